"""The per-step aggregates of --summary-output (MSNet::writeSummaryOutput) against the reference's own file, attribute by attribute
as the strings it prints.  The two mean speeds are sums of ~1e3 doubles added in another order than the reference's (id order
there, slot / tree order here): they may differ in the last printed digit when the mean sits on a rounding boundary, so they
are compared with a tolerance of 0.01 (stated here, the only one); everything else must be the identical string."""
import subprocess

import pytest

from conftest import reference_case, reference_summary, summary_strings
from sumo_b200.flatfile import load_trace

CASES = ["grid3_small", "grid4_tls", "grid5_dense", "grid4_tls_step05", "c1_grid10_depart0"]
FIELDS = ["time_ms", "inserted", "running", "waiting", "ended", "arrived", "collisions", "teleports", "halting", "stopped",
          "mean_waiting_time", "mean_travel_time", "mean_speed", "mean_speed_relative"]
TOLERANT = ("meanSpeed", "meanSpeedRelative")


def compare(ref_steps, got_steps):
    assert len(got_steps) >= 100
    seen_waiting = seen_ended = False
    for ref, got in zip(ref_steps, got_steps):
        for k, s in got.items():
            if k in TOLERANT:
                assert abs(float(ref[k]) - float(s)) <= 0.0100001, "t=%s: %s is %s, reference %s" % (ref["time"], k, s, ref[k])
            else:
                assert ref[k] == s, "t=%s: %s is %s, reference %s" % (ref["time"], k, s, ref[k])
        seen_waiting |= int(ref["waiting"]) > 0
        seen_ended |= int(ref["ended"]) > 0
    return seen_waiting, seen_ended


@pytest.mark.parametrize("case", CASES)
def test_host_emulation_summary_matches_reference(case, host_emul, tmp_path):
    flat, trace = reference_case(case)
    ref = reference_summary(case)
    n = len(load_trace(trace))
    out = str(tmp_path / "summary.txt")
    subprocess.run([host_emul, flat, "--summary-out", out, "--steps", str(n)], check=True, timeout=600, stderr=subprocess.DEVNULL)
    got = []
    for line in open(out):
        f = line.split()
        got.append(summary_strings({k: (float(x) if i >= 10 else int(x)) for i, (k, x) in enumerate(zip(FIELDS, f))}))
    assert len(got) == n
    _, ended = compare(ref, got)
    assert ended


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_engine_summary_matches_reference(case):
    from sumo_b200.engine import Engine
    flat, trace = reference_case(case)
    ref = reference_summary(case)
    n = min(len(load_trace(trace)), 1200)
    got = []
    with Engine(flat) as eng:
        for _ in range(n):
            r = eng.step(1)
            s = eng.summary()
            assert s["running"] == r.running and s["halting"] == r.halting and s["inserted"] == r.departed_total
            got.append(summary_strings(s))
        assert r.flags == 0
    compare(ref, got)
