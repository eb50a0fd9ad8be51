#!/bin/bash
# dev helper (runs ON the GPU box): the round's profile set -- bench line, ncu launch list, one full capture of k_plan and k_execute.
# Each ncu pass only after the plain command exited 0.  usage: profile.sh <tag>
tag=$1
timeout 600 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err || { echo bench failed; tail -5 gpurun_out/${tag}_bench.err; exit 1; }
tail -c 600 gpurun_out/${tag}_bench.json
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > /dev/null 2>&1 || { echo short bench failed; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 3384 -c 130 --csv --log-file gpurun_out/${tag}_launches.csv \
  python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_ncu1.log 2>&1; echo launches rc=$?
for k in k_plan k_execute; do
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$k --launch-skip 424 -c 1 -f -o gpurun_out/${tag}_$k \
  python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_ncu_$k.log 2>&1; echo $k rc=$?
done
ls -la gpurun_out | tail -8
