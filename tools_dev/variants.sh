#!/bin/bash
# dev helper: build tuning variants of the engine into /root/repo/build_variants/ (travels to the GPU box)
cd /root/repo; mkdir -p build_variants
F="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -prec-div=true -prec-sqrt=true -Xcompiler -fPIC,-ffp-contract=off -shared"
for v in "$@"; do
  p=${v%%:*}; x=${v##*:}
  nvcc $F -DSB2_PLAN_MINB=$p -DSB2_EXEC_MINB=$x -o build_variants/lib_p${p}_x${x}.so sumo_b200/csrc/engine.cu &
done
wait; ls -la build_variants
