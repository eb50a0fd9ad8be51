#!/bin/bash
# dev helper (1-GPU box): the round's validation -- GPU suite, default bench line, reference arm, C1 line
cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -6 | tee gpurun_out/r02_final_gpu_suite.txt
timeout 900 python bench.py > gpurun_out/r02_final_bench.json 2> gpurun_out/r02_final_bench.err; echo bench rc=$?
tail -c 600 gpurun_out/r02_final_bench.json
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_final_reference.json 2> gpurun_out/r02_final_reference.err; echo ref rc=$?
tail -c 500 gpurun_out/r02_final_reference.json
timeout 600 python tools_dev/c1_line.py > gpurun_out/r02_c1_line.txt 2> gpurun_out/r02_c1_line.err; echo c1 rc=$?
cat gpurun_out/r02_c1_line.txt; tail -3 gpurun_out/r02_c1_line.err
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
