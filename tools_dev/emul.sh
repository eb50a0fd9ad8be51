#!/bin/bash
# dev helper: build the host emulation into /tmp/w
set -e
mkdir -p /tmp/w
cd /root/repo/tests/host_emul && g++ -std=c++17 -O1 -g -ffp-contract=off -Wall -Wno-unused-variable -Wno-unused-but-set-variable -o /tmp/w/host_emul emul_main.cpp
