"""GPU suite: libsumob200.so through its C ABI (ctypes) against traces of the UNMODIFIED reference `sumo`.  Bit-exact."""
import numpy as np
import pytest

from conftest import reference_case
from sumo_b200.flatfile import load_flat, load_trace

pytestmark = pytest.mark.gpu

GPU_CASES = ["grid3_small", "grid3_small_depart0", "c1_grid10", "c1_grid10_depart0", "grid4_tls", "grid5_dense", "grid4_tls_speeddev", "grid4_tls_arrivalpos", "grid4_tls_mixed", "grid5_dense_mixed", "c1_grid10_mixed", "grid4_tls_step05", "grid5_dense_optmix", "rand_net", "grid4_stop", "grid3_tls_short",
             "grid4_tls_orig1", "grid5_dense_orig1", "grid4_tls_arrivalspeed"]


@pytest.mark.parametrize("case", GPU_CASES)
def test_engine_matches_reference(case):
    from sumo_b200.engine import Engine
    flat, trace = reference_case(case)
    ref = load_trace(trace)
    numid = load_flat(flat)["veh_numid"]
    order = np.argsort(numid, kind="stable")
    compared = 0
    with Engine(flat) as eng:
        for (t, ids, lanes, pos, speed) in ref:
            assert eng.time_ms == t
            r = eng.step(1)
            st = eng.vehicles()[order]
            on = st["lane"] >= 0
            assert r.flags == 0, "t=%d: engine raised flags 0x%x" % (t, r.flags)
            assert r.running - r.teleporting == int(on.sum()) == len(ids), "t=%d: running vehicles differ" % t
            assert r.halting == int((speed < 0.1).sum()), "t=%d: halting vehicles (summary output) differ" % t
            assert np.array_equal(numid[order][on], ids), "t=%d: sets of vehicles on the road differ" % t
            assert np.array_equal(st["lane"][on], lanes), "t=%d: lanes differ" % t
            assert np.array_equal(st["pos"][on].view(np.int64), pos.view(np.int64)), "t=%d: positions differ (bitwise)" % t
            assert np.array_equal(st["speed"][on].view(np.int64), speed.view(np.int64)), "t=%d: speeds differ (bitwise)" % t
            compared += len(ids)
        assert r.vehicle_updates == compared
    assert compared > 1000


@pytest.mark.parametrize("case,period", [("c1_grid10", 7), ("grid5_dense", 3), ("grid4_tls_mixed", 1), ("grid4_tls_hbefa4", 4)])
def test_reslotting_changes_nothing(case, period):
    """sumob200_set_reslot_period: the on-road vehicles are renumbered in lane order every few steps (memory locality); positions,
    speeds, lanes (and the emission sums) stay bit-identical to the reference, the API keeps speaking load-order indices"""
    from sumo_b200.engine import Engine
    flat, trace = reference_case(case)
    ref = load_trace(trace)
    numid = load_flat(flat)["veh_numid"]
    order = np.argsort(numid, kind="stable")
    compared = 0
    with Engine(flat) as eng:
        eng.set_reslot_period(period)
        for (t, ids, lanes, pos, speed) in ref:
            r = eng.step(1)
            st = eng.vehicles()[order]
            on = st["lane"] >= 0
            assert r.flags == 0, "t=%d: engine raised flags 0x%x" % (t, r.flags)
            assert np.array_equal(numid[order][on], ids), "t=%d: sets of vehicles on the road differ" % t
            assert np.array_equal(st["lane"][on], lanes), "t=%d: lanes differ" % t
            assert np.array_equal(st["pos"][on].view(np.int64), pos.view(np.int64)), "t=%d: positions differ (bitwise)" % t
            assert np.array_equal(st["speed"][on].view(np.int64), speed.view(np.int64)), "t=%d: speeds differ (bitwise)" % t
            compared += len(ids)
        if "hbefa" in case:
            from conftest import emissions_path, load_emissions
            te, eids, evals = load_emissions(emissions_path(case))[len(ref) - 1]
            got = eng.emissions()
            slot = {int(n): i for i, n in enumerate(numid)}
            idx = [slot[int(i)] for i in eids]
            assert np.array_equal(np.ascontiguousarray(got[:, idx].T).view(np.int64), evals.view(np.int64))
        rows = eng.fcd()
        assert sorted(rows["slot"].tolist()) == sorted(np.nonzero(eng.vehicles()["lane"] >= 0)[0].tolist())
    assert compared > 1000


def test_teleports_are_executed():
    """grid5_dense under the reference's default --time-to-teleport 300: the first vehicle is teleported at t = 484 s, 51 follow
    until t = 900 s (jam and yield teleports, re-insertions further along the route, teleports beyond the arrival edge); the
    trajectories stay bit-identical through all of them (test_engine_matches_reference[grid5_dense]) and the counter agrees with
    the reference's summary output"""
    from sumo_b200.engine import Engine
    from conftest import reference_summary
    flat, trace = reference_case("grid5_dense")
    ref = reference_summary("grid5_dense")
    with Engine(flat) as eng:
        r = eng.step(480)
        assert r.flags == 0 and r.teleports == int(ref[479]["teleports"]) == 0
        r = eng.step(420)
        assert r.flags == 0 and r.teleports == int(ref[899]["teleports"]) > 20


@pytest.mark.parametrize("case", ["grid4_tls_sigma", "grid5_dense_sigma", "c1_grid10_sigma", "bench_grid10_tls_draws"])
def test_engine_matches_reference_sigma_with_replayed_draws(case):
    """Krauss sigma=0.5 (BASELINE config 2's model): per-vehicle trajectories are bit-identical to the reference when the
    reference's per-vehicle RNG draws are replayed through sumob200_set_random_draws"""
    from sumo_b200.engine import Engine
    from conftest import draws_path, load_draws
    flat, trace = reference_case(case)
    ref = load_trace(trace)
    draws = load_draws(draws_path(case))
    assert all(ok for (_, _, _, ok) in draws)
    numid = load_flat(flat)["veh_numid"]
    slot = {int(n): i for i, n in enumerate(numid)}
    order = np.argsort(numid, kind="stable")
    compared = 0
    buf = np.zeros(len(numid))
    with Engine(flat) as eng:
        for (t, ids, lanes, pos, speed), (td, dids, dvals, _) in zip(ref, draws):
            assert td == t == eng.time_ms
            buf[[slot[int(i)] for i in dids]] = dvals
            eng.set_random_draws(buf)
            r = eng.step(1)
            st = eng.vehicles()[order]
            on = st["lane"] >= 0
            assert r.flags == 0, "t=%d: engine raised flags 0x%x" % (t, r.flags)
            assert np.array_equal(numid[order][on], ids), "t=%d: sets of vehicles on the road differ" % t
            assert np.array_equal(st["lane"][on], lanes), "t=%d: lanes differ" % t
            assert np.array_equal(st["pos"][on].view(np.int64), pos.view(np.int64)), "t=%d: positions differ (bitwise)" % t
            assert np.array_equal(st["speed"][on].view(np.int64), speed.view(np.int64)), "t=%d: speeds differ (bitwise)" % t
            compared += len(ids)
    assert compared > 1000


@pytest.mark.parametrize("case", ["grid3_small_idm", "grid5_dense_idm", "c1_grid10_idm", "grid4_tls_idm"])
def test_engine_matches_reference_idm(case):
    """carFollowModel IDM.  pow(x, delta) is not specified bit for bit by glibc (0.52 ulp) nor by CUDA, so this is the one
    floating-point tolerance of the suite: the same vehicles on the same lanes after every step, position and speed within
    1e-9 (m, m/s); the device computes x^4 correctly rounded in double-double arithmetic (measured deviation <= 6e-14)."""
    from sumo_b200.engine import Engine
    TOL = 1e-9
    flat, trace = reference_case(case)
    ref = load_trace(trace)
    numid = load_flat(flat)["veh_numid"]
    order = np.argsort(numid, kind="stable")
    compared = 0
    with Engine(flat) as eng:
        for (t, ids, lanes, pos, speed) in ref:
            r = eng.step(1)
            st = eng.vehicles()[order]
            on = st["lane"] >= 0
            assert r.flags == 0
            assert np.array_equal(numid[order][on], ids), "t=%d: sets of vehicles on the road differ" % t
            assert np.array_equal(st["lane"][on], lanes), "t=%d: lanes differ" % t
            assert np.abs(st["pos"][on] - pos).max(initial=0) <= TOL, "t=%d: positions differ by more than %g" % (t, TOL)
            assert np.abs(st["speed"][on] - speed).max(initial=0) <= TOL, "t=%d: speeds differ by more than %g" % (t, TOL)
            compared += len(ids)
    assert compared > 1000
