#!/bin/bash
# dev helper (1-GPU box): E2 detectors on the device + through the hook, then the whole GPU suite and the default bench line
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_e2_detectors.py tests/test_boundary.py -x -q -m gpu 2>&1 | tail -8 | tee gpurun_out/r02_e2_gpu.txt
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -6 | tee gpurun_out/r02_final_gpu_suite.txt
timeout 900 python bench.py > gpurun_out/r02_final_bench.json 2> gpurun_out/r02_final_bench.err; echo bench rc=$?
tail -c 400 gpurun_out/r02_final_bench.json
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
