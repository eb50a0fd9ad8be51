#!/bin/bash
# dev helper (1-GPU box): the allocation-free entry point's test alone in a fresh process, then the whole GPU suite (which collects
# it last), then the full-size c4 network on ONE unpartitioned engine with SUMOB200_DEBUG_SYNC (does the fault need partitions?)
cd /root/repo
mkdir -p gpurun_out
echo "== test_zz_no_alloc alone"
timeout 900 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -15 | tee gpurun_out/r02_noalloc_alone.txt
echo "== whole gpu suite"
timeout 2400 python -m pytest tests -x -q -m gpu 2>&1 | tail -15 | tee gpurun_out/r02_gpu_suite.txt
avail=$(awk '/MemAvailable/ {print int($2/1048576)}' /proc/meminfo)
echo "MemAvailable ${avail} GB"
if [ "$avail" -gt 110 ]; then
echo "== c4 full size, one engine, debug sync"
SUMOB200_DEBUG_SYNC=1 timeout 900 python tools_dev/c4_repro.py 1000 10000000 1 40 > gpurun_out/r02_c4_full1.txt 2>&1
echo rc=$?; tail -8 gpurun_out/r02_c4_full1.txt
fi
