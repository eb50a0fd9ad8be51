#!/bin/bash
# dev helper (1-GPU box): the whole GPU suite, the default bench line, smoke
cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -8 | tee gpurun_out/r02_final_gpu_suite.txt
timeout 900 python bench.py > gpurun_out/r02_final_bench.json 2> gpurun_out/r02_final_bench.err; echo bench rc=$?
tail -c 300 gpurun_out/r02_final_bench.json
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
