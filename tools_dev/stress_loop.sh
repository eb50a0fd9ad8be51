#!/bin/bash
# dev helper (GPU box): N spin-ups of the bench workload with the bounds-checked build; prints anything that is not the usual line
n=${1:-20}; lib=${2:-build_variants/lib_check.so}
for i in $(seq 1 $n); do SUMOB200_LIB=$PWD/$lib timeout 120 python tools_dev/stress.py 556 1 2>&1 | grep -v "^rep 0 flags 0 running 1000800 checksum 125668445.098545" | head -5; done; echo "stress loop of $n done"
