"""CPU suite: the step logic (device functions compiled for the host, tests/host_emul) against traces of the UNMODIFIED
reference `sumo` (oracle/_ref/sumo_flatten).  Bit-exact: lane index, fp64 position and fp64 speed of every vehicle on the
road after every step."""
import subprocess

import numpy as np
import pytest

from conftest import reference_case
from sumo_b200.flatfile import load_trace

HOST_CASES = ["grid3_small", "grid3_small_depart0", "c1_grid10", "grid4_tls", "grid5_dense", "grid4_tls_speeddev", "grid4_tls_arrivalpos", "grid4_tls_mixed", "grid5_dense_mixed", "c1_grid10_mixed", "grid4_tls_step05", "grid5_dense_optmix", "rand_net", "grid4_stop", "grid3_tls_short",
              # IDM: the host build calls glibc's pow like the reference, so it is bit-exact here (the GPU suite checks IDM with a tolerance)
              "grid3_small_idm", "grid5_dense_idm", "c1_grid10_idm", "grid4_tls_idm",
              "grid4_tls_orig1", "grid5_dense_orig1", "grid4_tls_arrivalspeed"]


def assert_traces_identical(ref, got):
    assert len(ref) == len(got), "number of steps differs: reference %d, engine %d" % (len(ref), len(got))
    n = 0
    for (ta, ida, la, pa, sa), (tb, idb, lb, pb, sb) in zip(ref, got):
        assert ta == tb
        assert np.array_equal(ida, idb), "t=%d: sets of vehicles on the road differ" % ta
        assert np.array_equal(la, lb), "t=%d: lanes differ" % ta
        assert np.array_equal(pa.view(np.int64), pb.view(np.int64)), "t=%d: positions differ (bitwise)" % ta
        assert np.array_equal(sa.view(np.int64), sb.view(np.int64)), "t=%d: speeds differ (bitwise)" % ta
        n += len(ida)
    return n


@pytest.mark.parametrize("case", HOST_CASES)
def test_host_emulation_matches_reference(case, host_emul, tmp_path):
    flat, trace = reference_case(case)
    out = str(tmp_path / "emul.trace")
    ref = load_trace(trace)
    subprocess.run([host_emul, flat, "--trace-out", out, "--steps", str(len(ref))], check=True, timeout=600, stderr=subprocess.DEVNULL)
    n = assert_traces_identical(ref, load_trace(out))
    assert n > 1000


SIGMA_CASES = ["grid4_tls_sigma", "grid5_dense_sigma", "c1_grid10_sigma", "bench_grid10_tls_draws"]


@pytest.mark.parametrize("case", SIGMA_CASES)
def test_host_emulation_matches_reference_with_replayed_draws(case, host_emul, tmp_path):
    """Krauss sigma=0.5: bit-exact when the reference's per-vehicle dawdle draws are replayed (BASELINE north_star)"""
    from conftest import draws_path, load_draws
    flat, trace = reference_case(case)
    draws = load_draws(draws_path(case))
    assert all(ok for (_, _, _, ok) in draws), "the harness' draw prediction was not confirmed by the reference's RNG call counters"
    out = str(tmp_path / "emul.trace")
    ref = load_trace(trace)
    subprocess.run([host_emul, flat, "--trace-out", out, "--draws", draws_path(case), "--steps", str(len(ref))], check=True, timeout=600,
                   stderr=subprocess.DEVNULL)
    n = assert_traces_identical(ref, load_trace(out))
    assert n > 1000
