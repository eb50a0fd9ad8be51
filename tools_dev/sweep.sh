#!/bin/bash
# dev helper (runs ON the GPU box): parity tests with the default lib, then bench every variant in build_variants/
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for f in build_variants/lib_*.so; do v=$(basename $f .so); SUMOB200_LIB=$PWD/$f timeout 180 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/$1_$v.log 2>&1; echo $v rc=$?; python -c "
import json,sys
l=json.loads(open('gpurun_out/$1_$v.log').read().strip().splitlines()[-1]); print('%.1f M/s'%(l['value']/1e6), {k:round(x,3) for k,x in l['roofline']['phase_ms_per_step'].items()})"; done
