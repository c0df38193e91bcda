#!/bin/bash
# One GPU call of the round: the no-alloc test alone, the whole GPU suite, smoke, the bench, leg removal, variants.
# usage (on the GPU box, from the repo root): bash tools/gpu_round.sh <tag>
tag=${1:-x}
o=gpurun_out
mkdir -p $o
python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > $o/zz_alone_$tag.log 2>&1; echo "alone rc=$?" >> $o/zz_alone_$tag.log
python -m pytest tests -x -q -m gpu > $o/gpu_tests_$tag.log 2>&1; echo "all rc=$?" >> $o/gpu_tests_$tag.log
python -c "import __graft_entry__ as g; g.smoke()" > $o/smoke_$tag.log 2>&1; echo "smoke rc=$?" >> $o/smoke_$tag.log
python bench.py > $o/bench_$tag.json 2> $o/bench_$tag.err
timeout 300 python tools/gpu_experiments.py 0 4 36 6 2 128 132 134 38 7 0 > $o/legs_v1_$tag.txt 2>&1
timeout 300 python tools/gpu_experiments.py --v2 0 4 36 6 2 128 132 134 38 7 0 > $o/legs_v2_$tag.txt 2>&1
timeout 600 python tools/gpu_variants.py > $o/variants_$tag.log 2>&1
tail -3 $o/zz_alone_$tag.log $o/gpu_tests_$tag.log $o/smoke_$tag.log
cat $o/legs_v1_$tag.txt $o/legs_v2_$tag.txt $o/variants_$tag.log
