#!/usr/bin/env python3
"""dev helper (GPU box): how far do IDM trajectories of the CUDA engine stay from the reference?"""
import sys, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..")); sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
from conftest import reference_case
from sumo_b200.engine import Engine
from sumo_b200.flatfile import load_flat, load_trace
for case in ["grid3_small_idm", "grid5_dense_idm", "c1_grid10_idm"]:
    flat, trace = reference_case(case)
    ref = load_trace(trace); numid = load_flat(flat)["veh_numid"]; order = np.argsort(numid, kind="stable")
    worst = 0.0; firstbit = None; firstlane = None; n = 0
    with Engine(flat) as eng:
        for k, (t, ids, lanes, pos, speed) in enumerate(ref):
            eng.step(1); st = eng.vehicles()[order]; on = st["lane"] >= 0
            if on.sum() != len(ids) or not np.array_equal(numid[order][on], ids) or not np.array_equal(st["lane"][on], lanes):
                firstlane = k; break
            d = max(np.abs(st["pos"][on] - pos).max(initial=0), np.abs(st["speed"][on] - speed).max(initial=0)); worst = max(worst, d)
            if firstbit is None and (not np.array_equal(st["pos"][on].view(np.int64), pos.view(np.int64))):
                firstbit = k
            n += len(ids)
    print(case, "steps", len(ref), "vehicle-steps", n, "first bit difference at step", firstbit, "first lane/set difference at step", firstlane, "max abs diff before that", worst)
