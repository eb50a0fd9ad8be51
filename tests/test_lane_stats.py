"""sumob200_lane_stats_all = libsumo::Lane::getLastStepVehicleNumber / HaltingNumber / MeanSpeed / Occupancy for every lane
(src/libsumo/Lane.cpp:209-272), checked against the REFERENCE's trajectories: the reference trace gives lane, position and speed of
every vehicle after every step, from which the lane getters are functions (MSLane::getVehicleNumber, getMeanSpeed: speeds added rear to
front, the speed limit for an empty lane; halting = speed < 0.1).  The netto occupancy also counts vehicles lapping in from the next
lanes; it is compared on the lanes where nothing laps in or out (all vehicles fully on the lane, nothing within two lanes downstream
close to its lane's start)."""
import numpy as np
import pytest

from conftest import reference_case
from sumo_b200.flatfile import load_flat, load_trace

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("case", ["grid4_tls", "grid5_dense"])
def test_lane_stats_match_the_reference_trajectories(case):
    from sumo_b200.engine import Engine
    flat, trace = reference_case(case)
    g = load_flat(flat)
    ref = load_trace(trace)
    nL = len(g["lane_length"])
    vlen = g["vt_length"][g["veh_type"]]
    slot_of = {int(n): i for i, n in enumerate(g["veh_numid"])}
    # lanes within two links downstream of a lane
    succ = [set() for _ in range(nL)]
    for k in range(len(g["link_from"])):
        f = int(g["link_from"][k])
        for t in (int(g["link_via"][k]), int(g["link_to"][k])):
            if t >= 0:
                succ[f].add(t)
    checked = occupied = occ_checked = 0
    with Engine(flat) as eng:
        for step, (t, ids, lanes, pos, speed) in enumerate(ref[:600]):
            eng.step(1, result=False)
            if step % 25 != 24:
                continue
            st = eng.lane_stats()
            count = np.bincount(lanes, minlength=nL)
            assert np.array_equal(st["vehicle_number"], count), "t=%d" % t
            assert np.array_equal(st["halting_number"], np.bincount(lanes[speed < 0.1], minlength=nL)), "t=%d" % t
            near_start = np.zeros(nL, bool)
            for lane, p, i in zip(lanes, pos, ids):
                if p < vlen[slot_of[int(i)]] + 1e-9:
                    near_start[lane] = True
            for lane in range(nL):
                k = np.flatnonzero(lanes == lane)
                if len(k) == 0:
                    assert st["mean_speed"][lane] == g["lane_speed"][lane]
                    continue
                k = k[np.argsort(pos[k], kind="stable")]
                v = 0.
                for s in speed[k]:
                    v += s
                assert st["mean_speed"][lane] == v / len(k), "t=%d lane %d" % (t, lane)
                occupied += 1
                down = set(succ[lane])
                for d in list(down):
                    down |= succ[d]
                for d in list(down):
                    down |= succ[d]
                if not near_start[lane] and not any(count[d] and near_start[d] for d in down):
                    netto = 0.
                    for i in ids[k]:
                        netto += vlen[slot_of[int(i)]]
                    assert st["occupancy"][lane] == netto / g["lane_length"][lane], "t=%d lane %d" % (t, lane)
                    occ_checked += 1
            assert (st["waiting_time"] >= 0).all() and (st["waiting_time"][count == 0] == 0).all()
            checked += 1
    assert checked >= 20 and occupied > 500 and occ_checked > 50
