#!/bin/bash
# on the GPU box: time every build/variants/v*.so with tools/gpu_experiments.py <flags...>
cp xtts_b200/libttts_mel.so /tmp/orig.so
for so in build/variants/v*.so; do
  echo "== $so: $(grep "^$(basename $so .so):" build/variants/list.txt)"
  cp $so xtts_b200/libttts_mel.so
  timeout 120 python tools/gpu_experiments.py "$@" 2>&1 | grep flags=
done
cp /tmp/orig.so xtts_b200/libttts_mel.so
