"""HBEFA emission device (MSDevice_Emissions + HelpersHBEFA3/4::compute): the 7 accumulators of every on-road vehicle after
every step, bitwise against the reference's own device objects (sumo_flatten --emissions-out)."""
import subprocess

import numpy as np
import pytest

from conftest import emissions_path, load_emissions, reference_case
from sumo_b200.flatfile import load_flat, load_trace

EMISSION_CASES = ["grid4_tls_hbefa4", "grid4_tls_hbefa3", "grid5_dense_hbefa3"]


@pytest.mark.parametrize("case", EMISSION_CASES)
def test_host_emulation_emissions_match_reference(case, host_emul, tmp_path):
    flat, trace = reference_case(case)
    ref = load_emissions(emissions_path(case))
    out = str(tmp_path / "emul.emis")
    subprocess.run([host_emul, flat, "--emissions-out", out, "--steps", str(len(ref))], check=True, timeout=600, stderr=subprocess.DEVNULL)
    got = load_emissions(out)
    assert len(ref) == len(got)
    n = 0
    for (ta, ia, va), (tb, ib, vb) in zip(ref, got):
        assert ta == tb and np.array_equal(ia, ib)
        assert np.array_equal(va.view(np.int64), vb.view(np.int64)), "t=%d: emission accumulators differ (bitwise)" % ta
        n += len(ia)
    assert n > 1000
    assert max(v[:, 0].max(initial=0) for (_, _, v) in ref) > 0   # CO2 really accumulates


@pytest.mark.gpu
@pytest.mark.parametrize("case", EMISSION_CASES)
def test_engine_emissions_match_reference(case):
    from sumo_b200.engine import Engine
    flat, trace = reference_case(case)
    ref = load_emissions(emissions_path(case))
    numid = load_flat(flat)["veh_numid"]
    slot = {int(n): i for i, n in enumerate(numid)}
    compared = 0
    with Engine(flat) as eng:
        for k, (t, ids, vals) in enumerate(ref):
            assert eng.time_ms == t
            r = eng.step(1)
            assert r.flags == 0
            if k % 7 == 0 or k == len(ref) - 1:   # the accumulators carry the whole history: sampling steps loses nothing
                em = eng.emissions()
                got = em[:, [slot[int(i)] for i in ids]].T
                assert np.array_equal(np.ascontiguousarray(got).view(np.int64), vals.view(np.int64)), "t=%d: emission accumulators differ (bitwise)" % t
                compared += len(ids)
    assert compared > 300
