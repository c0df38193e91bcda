#!/bin/bash
# usage: tools/variants.sh "<nvcc -D flags A>" "<flags B>" ... : builds one .so per flag set (CPU box, in parallel), to be
# timed on the GPU box by tools/gpu_variants.py
mkdir -p build/variants
rm -f build/variants/*.so build/variants/list.txt
i=0
for f in "$@"; do
  ( nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -shared -Xcompiler -fPIC -Iinclude $f -o build/variants/v$i.so xtts_b200/csrc/ttts_mel.cu || echo "v$i FAILED" ) &
  echo "v$i: $f" >> build/variants/list.txt
  i=$((i+1))
  if (( i % 6 == 0 )); then wait; fi
done
wait
ls build/variants/
