"""The committed golden vectors (tests/golden/grid3_small.golden.*, outputs of the unmodified reference): still what the
reference produces (CPU suite, where the oracle binary exists), and reproduced bit for bit by the CUDA engine (GPU suite)."""
import importlib.util
import os

import numpy as np
import pytest

from conftest import GOLDEN, REF
from sumo_b200.flatfile import load_flat, load_trace

FLAT = os.path.join(GOLDEN, "grid3_small.golden.flat")
TRACE = os.path.join(GOLDEN, "grid3_small.golden.trace")


def test_committed_golden_vectors_are_what_the_reference_produces(tmp_path):
    if not os.path.exists(os.path.join(REF, "sumo_flatten")):
        pytest.skip("oracle/_ref/sumo_flatten not built")
    spec = importlib.util.spec_from_file_location("make_golden_vectors", os.path.join(GOLDEN, "make_golden_vectors.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    flat, trace = str(tmp_path / "g.flat"), str(tmp_path / "g.trace")
    mod.generate(flat, trace)
    assert open(trace, "rb").read() == open(TRACE, "rb").read()
    a, b = load_flat(flat), load_flat(FLAT)
    assert sorted(a) == sorted(b)
    for k in a:
        assert np.array_equal(a[k], b[k]), k


def test_host_emulation_reproduces_golden_trace(host_emul, tmp_path):
    import subprocess
    ref = load_trace(TRACE)
    out = str(tmp_path / "e.trace")
    subprocess.run([host_emul, FLAT, "--trace-out", out, "--steps", str(len(ref))], check=True, timeout=300, stderr=subprocess.DEVNULL)
    assert open(out, "rb").read() == open(TRACE, "rb").read()


@pytest.mark.gpu
def test_engine_reproduces_golden_trace():
    from sumo_b200.engine import Engine
    ref = load_trace(TRACE)
    numid = load_flat(FLAT)["veh_numid"]
    order = np.argsort(numid, kind="stable")
    n = 0
    with Engine(FLAT) as eng:
        for (t, ids, lanes, pos, speed) in ref:
            r = eng.step(1)
            st = eng.vehicles()[order]
            on = st["lane"] >= 0
            assert r.flags == 0
            assert np.array_equal(numid[order][on], ids) and np.array_equal(st["lane"][on], lanes)
            assert np.array_equal(st["pos"][on].view(np.int64), pos.view(np.int64))
            assert np.array_equal(st["speed"][on].view(np.int64), speed.view(np.int64))
            n += len(ids)
    assert n == 3947
