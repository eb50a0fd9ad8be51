"""edgeData (src/microsim/output/MSMeanData.cpp, MSMeanData_Net.cpp) against the reference's own file.

The engine accumulates the RAW members of MSMeanData_Net::MSLaneMeanDataValues per lane (sumob200_meandata_interval); the derived
attributes of the <edge .../> line are the writer's (MSLaneMeanDataValues::write, restated below only to compare).  Reference:
`oracle/_ref/sumo -a <edgeData id file period="60"/>`.  Every attribute the reference prints must be there with the same value:
counters identical, real values within 0.01 of the printed two decimals (the per-lane sums are added with atomics, in another
order than the reference's lane-by-lane order)."""
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import CACHE, CASES, GOLDEN, REF, ensure_reference
from sumo_b200.flatfile import load_flat, load_trace

MD_CASES = ["grid4_tls", "grid5_dense", "c1_grid10", "grid4_tls_step05"]
PERIOD_S = 60
EPS = 0.001          # NUMERICAL_EPS
MAX_TT = 100000.     # maxTraveltime default of <edgeData>


def meandata_case(name):
    ensure_reference()
    os.makedirs(CACHE, exist_ok=True)
    base = name.replace("_step05", "")
    net, rou = os.path.join(GOLDEN, base + ".net.xml"), os.path.join(GOLDEN, base + ".rou.xml")
    add, out = os.path.join(CACHE, name + ".md.add.xml"), os.path.join(CACHE, name + ".md.out.xml")
    flat, trace = os.path.join(CACHE, name + ".md.flat"), os.path.join(CACHE, name + ".md.trace")
    exe = os.path.join(REF, "sumo_flatten")
    if not (os.path.exists(flat) and os.path.exists(out) and os.path.getmtime(flat) >= os.path.getmtime(exe)):
        with open(add, "w") as f:
            f.write('<additional>\n  <edgeData id="ed" file="%s" period="%d"/>\n</additional>\n' % (out, PERIOD_S))
        common = ["-n", net, "-r", rou, "-a", add, "--no-step-log", "--no-warnings"] + CASES[name]
        subprocess.run([os.path.join(REF, "sumo")] + common, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        os.replace(out, out + ".ref")
        subprocess.run([exe, "--flat-out", flat, "--trace-out", trace] + common, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        os.replace(out + ".ref", out)
    return flat, trace, out


def derive(r, lane_len, speed_limit, period=PERIOD_S):
    """MSLaneMeanDataValues::write (MSMeanData_Net.cpp:253-358) for one lane = one single-lane edge, default <edgeData> options"""
    sample, dist, wait, tloss, fsample, fdist, lensum, occ, minlen = (float(x) for x in r[:9])
    dep, arr, ent, left, tele = (int(x) for x in r[9:14])
    out = {"sampledSeconds": sample}
    cap = 1000. / max(minlen if minlen >= 0 else 1.7976931348623157e308, EPS)
    density = min(fsample / period * 1000. / lane_len, cap)
    overlap_density = min(sample / period * 1000. / lane_len, cap)
    if sample > 0:
        if fdist > EPS:
            out["traveltime"] = min(MAX_TT, lane_len * fsample / fdist)
        out["overlapTraveltime"] = min(MAX_TT, (lane_len + lensum / sample) * sample / dist) if dist > 0 else MAX_TT
        out["density"] = out["laneDensity"] = density
        out["occupancy"] = occ / period / lane_len * 100.
        out["waitingTime"], out["timeLoss"] = wait, tloss
        out["speed"] = dist / sample
        out["speedRelative"] = 0. if speed_limit == 0 else dist / sample / speed_limit
        out["flow"] = density * (fdist / fsample if fsample > 0 else 0.) * 3.6
        out["overlapDensity"] = overlap_density
    out.update(departed=dep, arrived=arr, entered=ent, left=left, laneChangedFrom=0, laneChangedTo=0)
    if tele:
        out["teleported"] = tele
    return out


def compare(ref_xml, got, g):
    """got: {(begin string, edge id): raw record (14 numbers)}"""
    nb, names = g["lane_name_begin"], bytes(g["lane_names"]).decode()
    lane_of_edge = {names[nb[i]:nb[i + 1]].rsplit("_", 1)[0]: i for i in range(len(nb) - 1) if not names[nb[i]:nb[i + 1]].startswith(":")}
    n = with_data = teleported = 0
    begin, full = None, False
    for line in open(ref_xml):
        m = re.search(r'<interval begin="([^"]*)" end="([^"]*)"', line)
        if m:
            begin, full = m.group(1), abs(float(m.group(2)) - float(m.group(1)) - PERIOD_S) < 1e-9
            continue
        m = re.search(r'<edge id="([^"]*)"(.*)/>', line)
        if not m or not full or (begin, m.group(1)) not in got:
            continue
        lane = lane_of_edge[m.group(1)]
        want = dict(re.findall(r' (\w+)="([^"]*)"', m.group(2)))
        mine = derive(got[(begin, m.group(1))], g["lane_length"][lane], g["lane_speed"][lane])
        assert set(want) == set(mine), (begin, m.group(1), want, mine)
        for k, v in want.items():
            if isinstance(mine[k], int):
                assert str(mine[k]) == v, (begin, m.group(1), k, v, mine[k])
            else:
                assert abs(float(v) - float("%.2f" % mine[k])) <= 0.0100001, (begin, m.group(1), k, v, mine[k])
        n += 1
        with_data += "traveltime" in want
        teleported += "teleported" in want
    assert n > 400 and with_data > 200
    return n, teleported


def lane_ids(g):
    nb, names = g["lane_name_begin"], bytes(g["lane_names"]).decode()
    return [names[nb[i]:nb[i + 1]] for i in range(len(nb) - 1)]


@pytest.mark.parametrize("case", MD_CASES)
def test_host_emulation_edgedata_matches_reference(case, host_emul, tmp_path):
    flat, trace, ref_xml = meandata_case(case)
    g = load_flat(flat)
    ids = lane_ids(g)
    out = str(tmp_path / "md.txt")
    subprocess.run([host_emul, flat, "--md-out", out, "--md-period", str(PERIOD_S * 1000), "--steps", str(len(load_trace(trace)))], check=True,
                   timeout=900, stderr=subprocess.DEVNULL)
    got = {}
    for line in open(out):
        f = line.split()
        got[("%.2f" % (int(f[0]) / 1000.), ids[int(f[1])].rsplit("_", 1)[0])] = f[2:]
    n, teleported = compare(ref_xml, got, g)
    if case == "grid5_dense":
        assert teleported > 10


@pytest.mark.gpu
@pytest.mark.parametrize("case", MD_CASES)
def test_engine_edgedata_matches_reference(case):
    from sumo_b200.engine import Engine
    flat, trace, ref_xml = meandata_case(case)
    g = load_flat(flat)
    ids = lane_ids(g)
    n = len(load_trace(trace))
    got = {}
    fields = ["sample_seconds", "travelled_distance", "wait_seconds", "time_loss", "front_sample_seconds", "front_travelled_distance", "veh_length_sum",
              "occupation_sum", "minimal_vehicle_length", "departed", "arrived", "entered", "left", "teleported"]
    with Engine(flat) as eng:
        eng.set_reslot_period(5)
        for _ in range(n):
            r = eng.step(1)
            assert r.flags == 0
            if r.time_ms % (PERIOD_S * 1000) == 0:
                md = eng.meandata_interval()
                for lane, rec in enumerate(md):
                    if not ids[lane].startswith(":"):
                        got[("%.2f" % (r.time_ms / 1000. - PERIOD_S), ids[lane].rsplit("_", 1)[0])] = [rec[k] for k in fields]
    compare(ref_xml, got, g)
