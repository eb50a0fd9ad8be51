#!/bin/bash
# dev helper: build tuning variants of the engine into /root/repo/build_variants/ (travels to the GPU box)
# usage: variants.sh name:"-Dflags" ...
cd /root/repo; mkdir -p build_variants
F="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -prec-div=true -prec-sqrt=true -Xcompiler -fPIC,-ffp-contract=off -shared"
for v in "$@"; do
  n=${v%%:*}; x=${v#*:}
  nvcc $F $x -o build_variants/lib_$n.so sumo_b200/csrc/engine.cu 2>/dev/null &
done
wait; ls build_variants
