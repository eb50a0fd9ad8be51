"""TraCI-style control between steps.  Traffic lights: sumob200_tls_set_phase against libsumo::TrafficLight::setPhase called on the UNMODIFIED
reference at the same times (the harness' --tls-set-phase).  29+ phase changes on a 4x4 traffic-light grid; the trajectories of
all vehicles stay bit-identical, so every signal state, its change time and the rescheduled switches are the reference's."""
import subprocess

import numpy as np
import pytest

from conftest import reference_case, tls_phase_events, veh_speed_events
from sumo_b200.flatfile import load_flat, load_trace, names
from test_parity_host import assert_traces_identical

CASE = "grid4_tls_setphase"


def events_by_index(flat):
    tl = {n: i for i, n in enumerate(names(load_flat(flat), "tl"))}
    return [(t, tl[tid], ph) for (t, tid, ph) in tls_phase_events("grid4_tls")]


def test_events_change_the_run():
    """the controlled run must differ from the uncontrolled one, else the comparison below proves nothing"""
    _, trace = reference_case(CASE)
    _, plain = reference_case("grid4_tls")
    a, b = load_trace(trace), load_trace(plain)
    assert any(len(x[1]) != len(y[1]) or not np.array_equal(x[3], y[3]) for x, y in zip(a, b))


def test_host_emulation_matches_reference_under_phase_control(host_emul, tmp_path):
    flat, trace = reference_case(CASE)
    ref = load_trace(trace)
    ev = events_by_index(flat)
    assert len(ev) >= 29
    out = str(tmp_path / "emul.trace")
    subprocess.run([host_emul, flat, "--trace-out", out, "--steps", str(len(ref)), "--tls-set-phase", ";".join("%d:%d:%d" % e for e in ev)],
                   check=True, timeout=600, stderr=subprocess.DEVNULL)
    assert assert_traces_identical(ref, load_trace(out)) > 50000


@pytest.mark.gpu
def test_engine_matches_reference_under_phase_control():
    from sumo_b200.engine import Engine, SumoB200Error
    flat, trace = reference_case(CASE)
    ref = load_trace(trace)
    numid = load_flat(flat)["veh_numid"]
    order = np.argsort(numid, kind="stable")
    pending = {}
    for (t, tl, ph) in events_by_index(flat):
        pending.setdefault(t, []).append((tl, ph))
    compared = 0
    with Engine(flat) as eng:
        assert eng.num_tls >= 4
        with pytest.raises(SumoB200Error, match="not in the allowed range"):
            eng.tls_set_phase(1, 99)
        for (t, ids, lanes, pos, speed) in ref:
            for (tl, ph) in pending.get(t, []):
                eng.tls_set_phase(tl, ph)
                got, next_switch = eng.tls_phase(tl)
                assert got == ph and next_switch > t
            r = eng.step(1)
            st = eng.vehicles()[order]
            on = st["lane"] >= 0
            assert r.flags == 0
            assert np.array_equal(numid[order][on], ids), "t=%d: sets of vehicles on the road differ" % t
            assert np.array_equal(st["lane"][on], lanes), "t=%d: lanes differ" % t
            assert np.array_equal(st["pos"][on].view(np.int64), pos.view(np.int64)), "t=%d: positions differ (bitwise)" % t
            assert np.array_equal(st["speed"][on].view(np.int64), speed.view(np.int64)), "t=%d: speeds differ (bitwise)" % t
            compared += len(ids)
    assert compared > 50000


# ---- libsumo::Vehicle::setSpeed ----------------------------------------------------------------------------------------------
SPEED_CASE = "grid4_tls_setspeed"


def speed_events_by_slot(flat):
    slot = {n: i for i, n in enumerate(names(load_flat(flat), "veh"))}
    return [(t, slot[vid], sp) for (t, vid, sp) in veh_speed_events("grid4_tls")]


def test_speed_advice_changes_the_run():
    _, trace = reference_case(SPEED_CASE)
    _, plain = reference_case("grid4_tls")
    a, b = load_trace(trace), load_trace(plain)
    differing = sum(1 for x, y in zip(a, b) if len(x[1]) != len(y[1]) or not np.array_equal(x[4], y[4]))
    assert differing > 500
    # advised speeds are really driven: some vehicle moves at exactly 3 m/s resp. 7.5 m/s at some time
    speeds = np.concatenate([x[4] for x in a])
    assert (speeds == 3.0).sum() > 20 and (speeds == 7.5).sum() > 20


def test_host_emulation_matches_reference_under_speed_advice(host_emul, tmp_path):
    flat, trace = reference_case(SPEED_CASE)
    ref = load_trace(trace)
    ev = speed_events_by_slot(flat)
    assert len(ev) > 350
    out = str(tmp_path / "emul.trace")
    subprocess.run([host_emul, flat, "--trace-out", out, "--steps", str(len(ref)), "--veh-set-speed", ";".join("%d:%d:%.17g" % e for e in ev)],
                   check=True, timeout=600, stderr=subprocess.DEVNULL)
    assert assert_traces_identical(ref, load_trace(out)) > 50000


@pytest.mark.gpu
def test_engine_matches_reference_under_speed_advice():
    from sumo_b200.engine import Engine, SumoB200Error
    flat, trace = reference_case(SPEED_CASE)
    ref = load_trace(trace)
    numid = load_flat(flat)["veh_numid"]
    order = np.argsort(numid, kind="stable")
    pending = {}
    for (t, slot, sp) in speed_events_by_slot(flat):
        pending.setdefault(t, []).append((slot, sp))
    compared = applied = 0
    with Engine(flat) as eng:
        with pytest.raises(SumoB200Error):
            eng.set_speed([eng.num_vehicles], [1.0])
        state = eng.vehicles()
        for (t, ids, lanes, pos, speed) in ref:
            ev = [(s, sp) for (s, sp) in pending.get(t, []) if state["lane"][s] >= 0]   # only vehicles on the road, like the harness
            if ev:
                eng.set_speed([s for s, _ in ev], [sp for _, sp in ev])
                applied += len(ev)
            r = eng.step(1)
            state = eng.vehicles()
            st = state[order]
            on = st["lane"] >= 0
            assert r.flags == 0
            assert np.array_equal(numid[order][on], ids), "t=%d: sets of vehicles on the road differ" % t
            assert np.array_equal(st["lane"][on], lanes), "t=%d: lanes differ" % t
            assert np.array_equal(st["pos"][on].view(np.int64), pos.view(np.int64)), "t=%d: positions differ (bitwise)" % t
            assert np.array_equal(st["speed"][on].view(np.int64), speed.view(np.int64)), "t=%d: speeds differ (bitwise)" % t
            compared += len(ids)
    assert compared > 50000 and applied > 200
