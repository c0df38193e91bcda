#!/bin/bash
# usage: tools/variants.sh "<nvcc -D flags A>" "<flags B>" ... : builds one .so per flag set (CPU box), to be timed by tools/gpu_variants.py
i=0
mkdir -p build/variants
rm -f build/variants/*.so
for f in "$@"; do
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -shared -Xcompiler -fPIC -Iinclude $f -o build/variants/v$i.so xtts_b200/csrc/ttts_mel.cu || exit 1
  echo "v$i: $f" >> build/variants/list.txt
  i=$((i+1))
done
