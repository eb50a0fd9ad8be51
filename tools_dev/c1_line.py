"""dev helper (GPU box): BASELINE configs[0] (C1 = tests/golden/c1_grid10: 10 x 10 grid, 2000 vehicles, the config that runs in CPU
sumo today), whole run: the reference binary (oracle/_ref/sumo, wall clock of the process incl. loading, and its own UPS line)
next to the engine through the C ABI (sumob200_step per step, device-timed; and with an FCD snapshot to the host every step).
Prints one JSON line; the trajectories of this case are compared bit by bit in tests/test_parity_gpu.py."""
import json
import os
import re
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import CASES, GOLDEN, REF, reference_case  # noqa: E402


def main():
    name = "c1_grid10"
    flat, _ = reference_case(name)
    cmd = [os.path.join(REF, "sumo"), "-n", os.path.join(GOLDEN, name + ".net.xml"), "-r", os.path.join(GOLDEN, name + ".rou.xml"),
           "--no-warnings", "--duration-log.statistics", "--no-step-log"] + CASES[name]
    t0 = time.time()
    out = subprocess.run(cmd, capture_output=True, text=True)
    ref_wall = time.time() - t0
    text = out.stdout + out.stderr
    ups = re.search(r"UPS:\s*([0-9.]+)", text)
    dur = re.search(r"Duration:\s*([0-9.]+)s", text)
    moved = re.search(r"Vehicles:.*?Inserted:\s*(\d+)", text, re.S)
    from sumo_b200.engine import Engine
    from sumo_b200.flatfile import load_flat
    g = load_flat(flat)
    eng = Engine(g, device=0)
    steps = 0
    t0 = time.time()
    ms_total = 0.
    r = None
    while True:
        r, ms = eng.step_timed(100)
        ms_total += ms
        steps += 100
        if (r.running == 0 and r.departed_total == eng.num_vehicles) or steps >= 3700:
            break
    wall = time.time() - t0
    updates = r.vehicle_updates
    eng.close()
    eng = Engine(g, device=0)
    t0 = time.time()
    rows = 0
    for _ in range(steps):
        eng.step(1, result=False)
        rows += len(eng.fcd())
    r2 = eng.step(0)
    e2e_wall = time.time() - t0
    eng.close()
    print(json.dumps({
        "config": "C1 c1_grid10 (2000 vehicles, whole run)",
        "reference": {"ups_reported": float(ups.group(1)) if ups else None, "duration_s_reported": float(dur.group(1)) if dur else None,
                      "process_wall_s": round(ref_wall, 3), "cores": 1, "inserted": int(moved.group(1)) if moved else None},
        "engine": {"steps": steps, "vehicle_updates": int(updates), "device_ms": round(ms_total, 2), "wall_s": round(wall, 3),
                   "ups_device": round(updates / (ms_total / 1e3)), "ups_wall": round(updates / wall),
                   "e2e_fcd_every_step": {"wall_s": round(e2e_wall, 3), "rows": rows, "ups": round(r2.vehicle_updates / e2e_wall)},
                   "flags": r.flags, "arrived": r.arrived_total},
        "note": "a network with ~150 vehicles on the road at a time cannot fill a GPU: the step is launch bound (12 launches, ~0.1 ms)"}))


if __name__ == "__main__":
    main()
