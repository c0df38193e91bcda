#!/bin/bash
# on the GPU box: time the resampler of every build/variants/v*.so (tools/gpu_resample_time.py <orig:new ...>)
cp xtts_b200/libttts_mel.so /tmp/orig.so
for so in build/variants/v*.so; do
  echo "== $so: $(grep "^$(basename $so .so):" build/variants/list.txt)"
  cp $so xtts_b200/libttts_mel.so
  timeout 120 python tools/gpu_resample_time.py "$@" 2>&1 | tail -1
done
cp /tmp/orig.so xtts_b200/libttts_mel.so
