#!/usr/bin/env python3
"""dev helper (GPU box): where does the end-to-end FCD streaming loop spend its time?"""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from sumo_b200.engine import Engine
from sumo_b200.flatfile import load_flat
from sumo_b200.tiler import tile
base = load_flat(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "bench_grid10_tls.flat"))
eng = Engine(tile(base, 556))
eng.step(430)
T = {"step": 0.0, "begin": 0.0, "end": 0.0}
def tm(k, f):
    t0 = time.perf_counter(); r = f(); T[k] += time.perf_counter() - t0; return r
n = 20
# serial: step(result) ; begin ; end
for _ in range(n):
    tm("step", lambda: eng.step(1)); tm("begin", eng.fcd_begin); tm("end", eng.fcd_end)
print("serial   ms/step:", {k: round(1000 * v / n, 3) for k, v in T.items()})
T = {k: 0.0 for k in T}
t0 = time.perf_counter()
eng.step(1, result=False); eng.fcd_begin()
for _ in range(n - 1):
    tm("step", lambda: eng.step(1, result=False)); tm("end", eng.fcd_end); tm("begin", eng.fcd_begin)
eng.fcd_end()
print("pipelined ms/step:", {k: round(1000 * v / n, 3) for k, v in T.items()}, "total", round(1000 * (time.perf_counter() - t0) / n, 3))
