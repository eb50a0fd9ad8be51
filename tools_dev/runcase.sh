#!/bin/bash
# dev helper: reference trace + host emulation trace + compare.  usage: runcase.sh <fixture> [extra sumo options...]
name=$1; shift
cd /tmp/w
/root/repo/oracle/_ref/sumo_flatten --flat-out $name.flat --trace-out $name.trace -n /root/repo/tests/golden/$name.net.xml -r /root/repo/tests/golden/$name.rou.xml --no-step-log --default.speeddev 0 "$@" > $name.ref.log 2>&1; cat $name.ref.log | grep -v SUMO_HOME | head -${WARN_LINES:-8}; true
timeout ${EMUL_TIMEOUT:-120} ./host_emul $name.flat --trace-out $name.e.trace ${EMUL_ARGS}
python3 /root/repo/tools_dev/cmp_trace.py $name.trace $name.e.trace $name.flat
