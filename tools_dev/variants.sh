#!/bin/bash
# dev helper: build tuning variants of the engine into /root/repo/build_variants/ (travels to the GPU box)
# usage: variants.sh name:"-Dflag -Dflag" ...
cd /root/repo; mkdir -p build_variants
for v in "$@"; do
  n=${v%%:*}; x=${v#*:}
  python - "$n" $x <<'PY'
import sys
sys.path.insert(0, "/root/repo")
from sumo_b200 import build
print(build.build(force=True, extra_flags=sys.argv[2:], out="/root/repo/build_variants/lib_%s.so" % sys.argv[1]))
PY
done
ls build_variants
