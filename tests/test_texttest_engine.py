"""The engine against the reference's OWN texttest goldens: FCD output (tests/sumo/output/fcd-output/*/fcd.sumo), E1 detectors
(tests/sumo/output/e1/**/detector.sumo), E2 lane area detectors (tests/sumo/output/e2/**/detector.sumo) and edge-based mean data (tests/sumo/output/meandata_edgebased/**/aggregated*.sumo) -- files
the reference's maintainers committed from a stock build: the test's inputs and options
are flattened by the reference's loader (oracle/_ref/sumo_flatten), the step runs in tests/host_emul (the device functions compiled
for the host), and the intervals it reports must be the committed lines -- every attribute as printed.

Needs /root/reference (the build container); skipped where it is absent."""
import os
import re
import shlex
import shutil
import subprocess

import pytest

from conftest import REF
from sumo_b200.flatfile import load_flat, load_trace

REFERENCE = os.environ.get("SUMO_REFERENCE", "/root/reference")
TT_ROOT = os.path.join(REFERENCE, "tests", "sumo")

E1_TESTS = ["output/e1/one_vehicle/" + t for t in ("one1s_1vehicle_10m_s", "one10s_1vehicle_10m_s", "one100s_1vehicle_10m_s", "two1s_1vehicle_10m_s",
                                                   "two10s_1vehicle_10m_s", "two100s_1vehicle_10m_s", "one1s_1vehicle_10m_shifted", "end_of_route")]
for group in ("two_same_vehicles", "two_vehicles_diff_length", "two_vehicles_diff_speed"):
    E1_TESTS += ["output/e1/%s/%s" % (group, t) for t in ("one1s_1vehicle_10m_s", "one10s_1vehicle_10m_s", "one100s_1vehicle_10m_s",
                                                             "two1s_1vehicle_10m_s", "two10s_1vehicle_10m_s", "two100s_1vehicle_10m_s")]


def stage(rel, tmp_path):
    """texttest semantics: options = concatenation of the options.sumo files from tests/sumo down to the test; inputs from the test
    directory and its ancestors (the nearest wins)"""
    parts = rel.split("/")
    dirs = [TT_ROOT] + [os.path.join(TT_ROOT, *parts[:i + 1]) for i in range(len(parts))]
    options = []
    for d in dirs:
        f = os.path.join(d, "options.sumo")
        if os.path.exists(f):
            options += shlex.split(open(f).read())
        for fn in os.listdir(d):
            p = os.path.join(d, fn)
            if os.path.isfile(p) and not fn.endswith(".sumo") and not fn.startswith("testsuite"):
                shutil.copy(p, str(tmp_path / fn))
    return options


def attrs(line):
    return dict(re.findall(r' (\w+)="([^"]*)"', line))


@pytest.mark.parametrize("rel", E1_TESTS)
def test_engine_reproduces_the_reference_e1_golden(rel, host_emul, tmp_path):
    if not os.path.isdir(os.path.join(TT_ROOT, rel)):
        pytest.skip("reference sources (tests/sumo) are not available on this machine")
    options = stage(rel, tmp_path)
    flat = str(tmp_path / "x.flat")
    r = subprocess.run([os.path.join(REF, "sumo_flatten"), "--flat-out", flat, "--trace-out", str(tmp_path / "x.trace")] + options,
                       cwd=str(tmp_path), capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout + r.stderr)[-2000:]
    add = open(str(tmp_path / "input_additional.add.xml")).read()
    freqs = set(re.findall(r'<e1Detector[^>]* (?:freq|period)="([^"]*)"', add))
    assert len(freqs) == 1
    period_ms = int(round(float(freqs.pop()) * 1000))
    g = load_flat(flat)
    nb, names = g["e1_name_begin"], bytes(g["e1_names"]).decode()
    ids = [names[nb[i]:nb[i + 1]] for i in range(len(nb) - 1)]
    out = str(tmp_path / "e1.txt")
    steps = len(load_trace(str(tmp_path / "x.trace")))     # the steps the reference ran (-e, or until the network was empty)
    subprocess.run([host_emul, flat, "--e1-out", out, "--e1-period", str(period_ms), "--steps", str(steps)], check=True, timeout=600,
                   stderr=subprocess.DEVNULL)
    got = {}
    for line in open(out):
        f = line.split()
        got[("%.2f" % (int(f[0]) / 1000.), ids[int(f[2])])] = (f[3],) + tuple("%.2f" % float(x) for x in f[4:9]) + (f[9],)
    golden = [attrs(l) for l in open(os.path.join(TT_ROOT, rel, "detector.sumo")) if "<interval" in l]
    assert len(golden) >= 1
    passages = 0
    for w in golden:
        key = (w["begin"], w["id"])
        assert key in got, key
        mine = got[key]
        want = (w["nVehContrib"], w["flow"], w["occupancy"], w["speed"], w["harmonicMeanSpeed"], w["length"], w["nVehEntered"])
        assert mine == want, (rel, key, mine, want)
        passages += int(w["nVehContrib"])
    assert passages >= 1 or any(int(w["nVehEntered"]) for w in golden)


E2_TESTS = ["output/e2/jam_test__preliminary/" + t for t in ("one1s", "one10s", "one100s")] + ["output/e2/one_vehicle/leave_early"]
for group, subsecond in (("one_vehicle", ""), ("one_vehicle_subsecond", ""), ("two_same_vehicles", ""), ("two_vehicles_diff_length", ""),
                         ("two_vehicles_diff_speed", "")):
    E2_TESTS += ["output/e2/%s/%s" % (group, t) for t in ("one1s_1vehicle_10m_s", "one10s_1vehicle_10m_s", "one100s_1vehicle_10m_s",
                                                           "two1s_1vehicle_10m_s", "two10s_1vehicle_10m_s", "two100s_1vehicle_10m_s")]
E2_TESTS.append("output/e2/one_vehicle_subsecond/one.1s_1vehicle_10m_s")


@pytest.mark.parametrize("rel", E2_TESTS)
def test_engine_reproduces_the_reference_e2_golden(rel, host_emul, tmp_path):
    """all 23 attributes of every <interval> of the committed lane area detector output (jams, halting durations, occupancy, ...)"""
    if not os.path.isdir(os.path.join(TT_ROOT, rel)):
        pytest.skip("reference sources (tests/sumo) are not available on this machine")
    import test_e2_detectors
    options = stage(rel, tmp_path)
    flat = str(tmp_path / "x.flat")
    r = subprocess.run([os.path.join(REF, "sumo_flatten"), "--flat-out", flat, "--trace-out", str(tmp_path / "x.trace"),
                        "--draws-out", str(tmp_path / "x.draws")] + options, cwd=str(tmp_path), capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout + r.stderr)[-2000:]
    add = open(str(tmp_path / "input_additional.add.xml")).read()
    freqs = set(re.findall(r'<(?:e2Detector|laneAreaDetector)[^>]* (?:freq|period)="([^"]*)"', add))
    assert len(freqs) == 1
    period_ms = int(round(float(freqs.pop()) * 1000))
    g = load_flat(flat)
    nb, names = g["e2_name_begin"], bytes(g["e2_names"]).decode()
    ids = [names[nb[i]:nb[i + 1]] for i in range(len(nb) - 1)]
    out = str(tmp_path / "e2.txt")
    steps = len(load_trace(str(tmp_path / "x.trace")))
    subprocess.run([host_emul, flat, "--e2-out", out, "--e2-period", str(period_ms), "--draws", str(tmp_path / "x.draws"), "--steps", str(steps)],
                   check=True, timeout=600, stderr=subprocess.DEVNULL)
    got = {}
    for line in open(out):
        f = line.split()
        got[("%.2f" % (int(f[0]) / 1000.), ids[int(f[2])])] = {a: test_e2_detectors.fmt(a, float(x)) for a, x in zip(test_e2_detectors.ATTRS, f[3:])}
    golden = [attrs(l) for l in open(os.path.join(TT_ROOT, rel, "detector.sumo")) if "<interval" in l]
    assert len(golden) >= 1
    seen = 0
    for w in golden:
        key = (w["begin"], w["id"])
        assert key in got, key
        for a in test_e2_detectors.ATTRS:
            # (the committed files print the count startedHalts as "3.00", this tree's writer as "3": counts are compared as numbers)
            same = float(got[key][a]) == float(w[a]) if a in test_e2_detectors.INTEGER else got[key][a] == w[a]
            assert same, (rel, key, a, got[key][a], w[a])
        seen += int(w["nVehEntered"])
    assert seen >= 1


# (test directory, committed golden, period in s)
MD_TESTS = [
    ("output/meandata_edgebased/one_vehicle/interval1s", "aggregated1.sumo", 1),
    ("output/meandata_edgebased/one_vehicle/interval15s", "aggregated15.sumo", 15),
    ("output/meandata_edgebased/one_vehicle/interval60s", "aggregated60.sumo", 60),
    ("output/meandata_edgebased/one_vehicle/interval15s60s", "aggregated15.sumo", 15),
    ("output/meandata_edgebased/one_vehicle/interval15s60s", "aggregated60.sumo", 60),
    ("output/meandata_edgebased/jam1", "aggregated15.sumo", 15),
    ("output/meandata_edgebased/jam1", "aggregated60.sumo", 60),
    ("output/meandata_edgebased/jam2", "aggregated15.sumo", 15),
    ("output/meandata_edgebased/jam2", "aggregated60.sumo", 60),
    ("output/meandata_edgebased/short_edge", "aggregated15.sumo", 15),
]


# tests of the same suite whose scenario uses something the engine refuses (message of integration/sumo_flatten.cpp)
MD_REFUSED = {
    "output/meandata_edgebased/jam1": "variableSpeedSigns are not supported",
    "output/meandata_edgebased/jam2": "variableSpeedSigns are not supported",
    "output/meandata_edgebased/short_edge": "<stop>s are not supported",
}


@pytest.mark.parametrize("rel,golden,period", MD_TESTS)
def test_engine_reproduces_the_reference_edgedata_golden(rel, golden, period, host_emul, tmp_path):
    """the raw per-lane sums of the engine, pushed through MSLaneMeanDataValues::write as restated in tests/test_meandata.py:derive,
    against the committed <edge .../> lines (excludeEmpty: only edges with data are in the golden; begin / end of a definition only
    select which intervals are written)"""
    if not os.path.isdir(os.path.join(TT_ROOT, rel)):
        pytest.skip("reference sources (tests/sumo) are not available on this machine")
    import test_meandata
    options = stage(rel, tmp_path)
    flat = str(tmp_path / "x.flat")
    # (the vehicles of these tests use DEFAULT_VEHTYPE, sigma 0.5: the reference's dawdle draws are recorded and replayed)
    r = subprocess.run([os.path.join(REF, "sumo_flatten"), "--flat-out", flat, "--trace-out", str(tmp_path / "x.trace"),
                        "--draws-out", str(tmp_path / "x.draws")] + options, cwd=str(tmp_path), capture_output=True, text=True, timeout=900)
    if rel in MD_REFUSED:
        assert r.returncode != 0 and MD_REFUSED[rel] in r.stderr, (r.stdout + r.stderr)[-2000:]      # outside the engine's scope: refused, not simplified
        return
    assert r.returncode == 0, (r.stdout + r.stderr)[-2000:]
    g = load_flat(flat)
    nb, names = g["lane_name_begin"], bytes(g["lane_names"]).decode()
    lane_ids = [names[nb[i]:nb[i + 1]] for i in range(len(nb) - 1)]
    if max(np_count(g["edge_lane_begin"])) > 1:
        pytest.skip("multi-lane network")
    lane_of_edge = {n.rsplit("_", 1)[0]: i for i, n in enumerate(lane_ids) if not n.startswith(":")}
    out = str(tmp_path / "md.txt")
    steps = len(load_trace(str(tmp_path / "x.trace")))
    subprocess.run([host_emul, flat, "--md-out", out, "--md-period", str(period * 1000), "--steps", str(steps), "--draws", str(tmp_path / "x.draws")], check=True, timeout=900,
                   stderr=subprocess.DEVNULL)
    got = {}
    for line in open(out):
        f = line.split()
        got[("%.2f" % (int(f[0]) / 1000.), lane_ids[int(f[1])].rsplit("_", 1)[0])] = f[2:]
    compared = 0
    begin = None
    for line in open(os.path.join(TT_ROOT, rel, golden)):
        m = re.search(r'<interval begin="([^"]*)" end="([^"]*)"', line)
        if m:
            begin = m.group(1)
            length = float(m.group(2)) - float(m.group(1))     # (the last interval of a run may be a partial one)
            continue
        m = re.search(r'<edge id="([^"]*)"(.*)/>', line)
        if not m:
            continue
        want = attrs(m.group(2))
        lane = lane_of_edge[m.group(1)]
        mine = test_meandata.derive(got[(begin, m.group(1))], g["lane_length"][lane], g["lane_speed"][lane], length)
        assert set(want) == set(mine), (begin, m.group(1), want, mine)
        for k, v in want.items():
            if isinstance(mine[k], int):
                assert str(mine[k]) == v, (begin, m.group(1), k, v, mine[k])
            else:
                assert "%.2f" % mine[k] == v, (begin, m.group(1), k, v, mine[k])
        compared += 1
    assert compared >= 3


def np_count(begin):
    return [int(begin[i + 1]) - int(begin[i]) for i in range(len(begin) - 1)]


FCD_TESTS = ["begin", "period", "period_begin", "subsecond", "precision", "angle"]


@pytest.mark.parametrize("name", FCD_TESTS)
def test_engine_reproduces_the_reference_fcd_golden(name, host_emul, tmp_path):
    """every <timestep> of the committed fcd.sumo (the device.fcd.begin / period options only select which steps are written): the
    <vehicle .../> lines of the engine's formatter (sb2_fcdxml.h) for that step must be the same strings -- positions, speeds and
    angles of the engine's trajectories printed like MSFCDExport does, incl. 5 ms and 25 ms step lengths (3 decimals)"""
    rel = "output/fcd-output/" + name
    if not os.path.isdir(os.path.join(TT_ROOT, rel)):
        pytest.skip("reference sources (tests/sumo) are not available on this machine")
    options = stage(rel, tmp_path)
    flat = str(tmp_path / "x.flat")
    r = subprocess.run([os.path.join(REF, "sumo_flatten"), "--flat-out", flat, "--trace-out", str(tmp_path / "x.trace"),
                        "--draws-out", str(tmp_path / "x.draws")] + options, cwd=str(tmp_path), capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, (r.stdout + r.stderr)[-2000:]
    steps = len(load_trace(str(tmp_path / "x.trace")))
    out = str(tmp_path / "engine.xml")
    subprocess.run([host_emul, flat, "--fcd-xml-out", out, "--draws", str(tmp_path / "x.draws"), "--steps", str(steps)], check=True, timeout=900,
                   stderr=subprocess.DEVNULL)

    def blocks(path):
        res, cur = {}, None
        for line in open(path):
            m = re.search(r'<timestep time="([^"]*)"', line)
            if m:
                cur = m.group(1)
                res[cur] = []
            elif "<vehicle" in line and cur is not None:
                res[cur].append(line.strip())
        return res

    golden, mine = blocks(os.path.join(TT_ROOT, rel, "fcd.sumo")), blocks(out)
    assert len(golden) >= 1 and sum(len(v) for v in golden.values()) >= 1
    for t, lines in golden.items():
        assert mine.get(t) == lines, (t, lines[:2], (mine.get(t) or [])[:2])
