"""dev helper (1-GPU box): what the row-f4 outputs cost at scale.  bench.py's default workload (connected 100x100 grid, 1 M vehicles) with
an E1 loop in the middle of every normal lane, an E2 detector over the last 100 m of every normal lane and edgeData on: 60 steps timed
on the device with the detectors on (kernel set `krauss`, + k_e2_update every step), then the three interval read-outs a 60 s period
would do (sumob200_e1_interval / _e2_interval / _meandata_interval, host wall time), next to the same 60 steps without detectors.
One JSON line."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def add_observers(g):
    normal = np.flatnonzero((g["lane_flags"] & 1) == 0).astype(np.int32)     # LF_INTERNAL = 1
    length = g["lane_length"][normal]
    g["e1_lane"] = normal
    g["e1_pos"] = length / 2
    g["e1_length"] = np.zeros(len(normal))
    g["e2_lane"] = normal
    g["e2_start"] = np.maximum(length - 100., 0.)
    g["e2_length"] = length - g["e2_start"]
    g["e2_jam_time"] = np.full(len(normal), 1000, np.int64)
    g["e2_jam_speed"] = np.full(len(normal), 5 / 3.6)
    g["e2_jam_dist"] = np.full(len(normal), 10.)
    g["md_enabled"] = np.array([1], np.int32)
    return len(normal)


def run(with_observers):
    from sumo_b200.engine import Engine
    w = bench.use_workload("c2")
    g = bench.build_scenario(w["grid"], w["vehicles"], w["sigma"])
    n_det = add_observers(g) if with_observers else 0
    eng = Engine(g, device=0)
    eng.set_reslot_period(64)
    del g
    t = 0
    while t < bench.SPINUP_STEPS:
        eng.step(60, result=False)
        t += 60
        if with_observers:
            eng.e1_interval(eng.time_ms - 60000, eng.time_ms)
            eng.e2_interval()
            eng.meandata_interval()
    r0 = eng.step(5)
    r1, ms = eng.step_timed(60)
    out = {"steps": 60, "device_ms_per_step": ms / 60, "updates_per_s": (r1.vehicle_updates - r0.vehicle_updates) / (ms / 1000.), "flags": int(r1.flags)}
    if with_observers:
        t0 = time.perf_counter()
        e1 = eng.e1_interval(eng.time_ms - 60000, eng.time_ms)
        t1 = time.perf_counter()
        e2 = eng.e2_interval()
        t2 = time.perf_counter()
        md = eng.meandata_interval()
        t3 = time.perf_counter()
        out.update({"detectors_each": n_det, "e1_interval_s": t1 - t0, "e2_interval_s": t2 - t1, "meandata_interval_s": t3 - t2,
                    "e1_passages_in_60_steps": int(e1["n_veh_contrib"].sum()), "e2_sampled_seconds": float(e2["sampled_seconds"].sum()),
                    "e2_jam_vehicle_steps": int(e2["jam_length_in_vehicles_sum"].sum()), "edgedata_sample_seconds": float(md["sample_seconds"].sum())})
    eng.close()
    return out


def main():
    on = run(True)
    off = run(False)
    print(json.dumps({"workload": "bench.py c2 (connected 100x100 grid, 1 M vehicles, Krauss sigma 0.5) + E1 loop, E2 detector (last 100 m) and edgeData on every normal lane",
                      "with_observers": on, "without": off, "step_cost_ratio": on["device_ms_per_step"] / off["device_ms_per_step"],
                      "note": "e1/e2/meandata_interval_s = host wall time of ONE read-out of all detectors (what a 60 s period costs once per 60 steps)"}))


if __name__ == "__main__":
    main()
