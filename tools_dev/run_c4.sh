#!/bin/bash
# dev helper (8-GPU box): BASELINE configs[3] -- 10 M vehicles on the 1000x1000 / 100 m grid across 8 GPUs -- at the largest size the
# host memory safely allows (every rank builds the whole scenario: ~72 GB peak per rank at full size)
cd /root/repo
avail=$(awk '/MemAvailable/ {print int($2/1048576)}' /proc/meminfo)
echo "MemAvailable ${avail} GB, $(nproc) cores"
if [ "$avail" -gt 750 ]; then G=1000; V=10000000; elif [ "$avail" -gt 400 ]; then G=700; V=4900000; else G=500; V=2500000; fi
echo "running grid $G vehicles $V"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29751 bench.py --config c4 --grid $G --vehicles $V --gpus 8 --steps 30 --warmup 5 > gpurun_out/r02_c4_8gpu.json 2> gpurun_out/r02_c4_8gpu.err
echo rc=$?
tail -c 300 gpurun_out/r02_c4_8gpu.err
