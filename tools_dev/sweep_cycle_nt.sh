#!/bin/bash
# dev helper (runs ON the GPU box): one-cycle bench for every variant in build_variants/ (no tests)
for f in build_variants/lib_*.so; do v=$(basename $f .so); SUMOB200_LIB=$PWD/$f timeout 180 python bench.py --no-cpu-baseline > gpurun_out/$1_$v.log 2>&1; python -c "
import json,sys
l=json.loads(open('gpurun_out/$1_$v.log').read().strip().splitlines()[-1]); print('$v', '%.1f M/s'%(l['value']/1e6), 'e2e %.1f'%(l['e2e']['value']/1e6), {k:round(x,3) for k,x in l['roofline']['phase_ms_per_step'].items()}, l['config']['engine_flags'])"; done
