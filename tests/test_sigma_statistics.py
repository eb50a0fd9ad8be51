"""sigma > 0 WITHOUT replay: the engine's own counter-based generator (sumob200_set_random_mode 1, the default when any vType has
sigma > 0) draws other numbers than the reference's mt19937 lane streams, so trajectories differ; north_star asks that the
aggregate tripinfo statistics agree within a stated tolerance.  Stated here: over the vehicles that arrived in both runs, mean
duration and mean route length within 2 %, mean timeLoss within 10 %, mean waitingTime within 15 % (or 2 s), arrived count within
2 %.  (The bit-exact sigma > 0 comparison with replayed draws is tests/test_parity_*.py.)"""
import subprocess

import numpy as np
import pytest

from conftest import reference_case, reference_tripinfo
from sumo_b200.flatfile import load_flat, load_trace, names

# an unsaturated network (BASELINE configs[0] demand with sigma 0.5).  Saturated ones are no basis for a tolerance: grid4_tls_sigma
# (800 vehicles on a 4x4 traffic-light grid, gridlock-prone) ends with 446 arrivals in the reference at its default seed and
# with 523 ... 749 at --seed 1..6, i.e. the reference disagrees with itself by +-20 %; the engine's own generator gives 633.
CASES = ["c1_grid10_sigma"]


def check(ref, got):
    """ref: {id: attribute strings}; got: {id: (duration s, routeLength m, timeLoss s, waitingTime s)}"""
    assert abs(len(got) - len(ref)) <= 0.02 * len(ref), (len(got), len(ref))
    common = sorted(set(ref) & set(got))
    assert len(common) >= 0.97 * len(ref) and len(common) > 300
    r = np.array([[float(ref[v][k]) for k in ("duration", "routeLength", "timeLoss", "waitingTime")] for v in common]).mean(axis=0)
    g = np.array([got[v] for v in common]).mean(axis=0)
    assert abs(g[0] - r[0]) <= 0.02 * r[0], ("mean duration", g[0], r[0])
    assert abs(g[1] - r[1]) <= 0.02 * r[1], ("mean routeLength", g[1], r[1])
    assert abs(g[2] - r[2]) <= 0.10 * r[2], ("mean timeLoss", g[2], r[2])
    assert abs(g[3] - r[3]) <= max(0.15 * r[3], 2.0), ("mean waitingTime", g[3], r[3])
    return r, g


@pytest.mark.parametrize("case", CASES)
def test_host_emulation_own_generator_statistics(case, host_emul, tmp_path):
    flat, trace = reference_case(case)
    ref = reference_tripinfo(case)
    vn = names(load_flat(flat), "veh")
    out = str(tmp_path / "trip.txt")
    subprocess.run([host_emul, flat, "--tripinfo-out", out, "--steps", str(len(load_trace(trace)))], check=True, timeout=600, stderr=subprocess.DEVNULL)
    got = {}
    for line in open(out):
        f = line.split()
        # slot depart departLane departPos departSpeed departDelay arrival arrivalLane arrivalPos arrivalSpeed duration routeLength waitingTime waitingCount timeLoss
        got[vn[int(f[0])]] = (int(f[10]) / 1000., float(f[11]), int(f[14]) / 1000., int(f[12]) / 1000.)
    check(ref, got)


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_engine_own_generator_statistics(case):
    from sumo_b200.engine import Engine
    flat, trace = reference_case(case)
    ref = reference_tripinfo(case)
    vn = names(load_flat(flat), "veh")
    with Engine(flat) as eng:
        eng.set_random_mode(1)
        r = eng.step(len(load_trace(trace)))
        assert r.flags == 0
        trips = eng.tripinfo()
    got = {vn[i]: (t["duration_ms"] / 1000., t["route_length"], t["time_loss_ms"] / 1000., t["waiting_ms"] / 1000.)
           for i, t in enumerate(trips) if t["arrival_ms"] >= 0}
    r_, g_ = check(ref, got)
    print(case, "reference means", r_, "engine means", g_)
