#!/usr/bin/env python3
"""dev helper (GPU box): spin-up + some steps of the bench workload, prints flags; exit code != 0 on a CUDA fault"""
import sys, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from sumo_b200.engine import Engine
from sumo_b200.flatfile import load_flat
from sumo_b200.tiler import tile
base = load_flat(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "bench_grid10_tls.flat"))
tiles = int(sys.argv[1]) if len(sys.argv) > 1 else 556
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 1
sc = tile(base, tiles)
for r in range(reps):
    eng = Engine(sc)
    eng.step(420, result=False)
    res = eng.step(60)
    st = eng.vehicles()
    print("rep", r, "flags", res.flags, "running", res.running, "checksum %.6f" % float(st["pos"][st["lane"] >= 0].sum()), flush=True)
    eng.close()
