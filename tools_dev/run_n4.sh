#!/bin/bash
# dev helper (2-GPU box): the driver's N=2 bench line (weak scaling: one connected network of 2 M vehicles in two strips)
cd /root/repo
mkdir -p gpurun_out
timeout 800 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29754 bench.py --gpus 4 --steps 20 --warmup 3 > gpurun_out/r02_final_n4.json 2> gpurun_out/r02_final_n4.err
echo rc=$?
tail -c 700 gpurun_out/r02_final_n4.json; tail -c 300 gpurun_out/r02_final_n4.err
