"""Per-trip records (MSDevice_Tripinfo::generateOutput) of the arrived vehicles against the reference's own --tripinfo-output,
attribute by attribute at its output precision (2 decimals)."""
import subprocess

import numpy as np
import pytest

from conftest import reference_case, reference_tripinfo, tripinfo_strings
from sumo_b200.flatfile import load_flat, load_trace, names

CASES = ["grid3_small", "grid4_tls", "c1_grid10"]
FIELDS = ["depart_ms", "depart_lane", "depart_pos", "depart_speed", "depart_delay_ms", "arrival_ms", "arrival_lane", "arrival_pos", "arrival_speed",
          "duration_ms", "route_length", "waiting_ms", "waiting_count", "time_loss_ms", "speed_factor"]


def compare(ref, got_by_name):
    n = 0
    for vid, rec in got_by_name.items():
        assert vid in ref, "engine reports an arrival of %s the reference does not" % vid
        for k, s in rec.items():
            assert ref[vid][k] == s, "tripinfo of %s: %s is %s, reference %s" % (vid, k, s, ref[vid][k])
        n += 1
    return n


@pytest.mark.parametrize("case", CASES)
def test_host_emulation_tripinfo_matches_reference(case, host_emul, tmp_path):
    flat, trace = reference_case(case)
    ref = reference_tripinfo(case)
    arrays = load_flat(flat)
    vn, ln = names(arrays, "veh"), names(arrays, "lane")
    out = str(tmp_path / "trip.txt")
    subprocess.run([host_emul, flat, "--tripinfo-out", out, "--steps", str(len(load_trace(trace)))], check=True, timeout=600, stderr=subprocess.DEVNULL)
    got = {}
    for line in open(out):
        f = line.split()
        rec = {k: (float(x) if "." in x or "e" in x or "inf" in x else int(x)) for k, x in zip(FIELDS, f[1:])}
        got[vn[int(f[0])]] = tripinfo_strings(rec, ln)
    assert compare(ref, got) == len(ref) > 20


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_engine_tripinfo_matches_reference(case):
    from sumo_b200.engine import Engine
    flat, trace = reference_case(case)
    ref = reference_tripinfo(case)
    arrays = load_flat(flat)
    vn, ln = names(arrays, "veh"), names(arrays, "lane")
    with Engine(flat) as eng:
        r = eng.step(len(load_trace(trace)))
        assert r.flags == 0
        trips = eng.tripinfo()
    got = {vn[i]: tripinfo_strings(trips[i], ln) for i in range(len(trips)) if trips[i]["arrival_ms"] >= 0}
    assert compare(ref, got) == len(ref) > 20
