#!/bin/bash
# dev helper (GPU box): bench the default build and the tuning variants in build_variants/ on the 100x100 / 1M workload
cd /root/repo
run() { SUMOB200_LIB=$2 python bench.py --steps 30 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$1', 'value %.1fM ms %.3f e2e %.1fM' % (d['value']/1e6, d['ms_per_step'], d['e2e']['value']/1e6), {k: round(v,3) for k,v in d['roofline']['phase_ms_per_step'].items()}, d['config']['engine_flags'])
"; }
run default ""
for f in build_variants/lib_*.so; do run $(basename $f .so) /root/repo/$f; done
