#!/bin/bash
# 1/2/4/8-GPU bench lines + config-4 sweep on one box (run under `gpurun --gpus 8`); results in gpurun_out/scale_*.json
mkdir -p gpurun_out
python bench.py --no-cpu-baseline > gpurun_out/scale_bench_1.json 2>/dev/null
python tools/sweep_bench.py > gpurun_out/scale_sweep_1.json 2>/dev/null
for n in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n > gpurun_out/scale_bench_$n.json 2>gpurun_out/scale_bench_$n.err
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 tools/sweep_bench.py > gpurun_out/scale_sweep_$n.json 2>gpurun_out/scale_sweep_$n.err
done
for n in 1 2 4 8; do
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/scale_bench_$n.json").read().strip().splitlines()[-1])
    print("bench N=$n", round(d["value"]), "audio-s/s", round(d["ms_per_step"], 4), "ms/step", d["clocks"])
except Exception as e:
    print("bench N=$n failed", e)
try:
    print("sweep N=$n", open("gpurun_out/scale_sweep_$n.json").read().strip().splitlines()[-1][:400])
except Exception as e:
    print("sweep N=$n failed", e)
PY
done
