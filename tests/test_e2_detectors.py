"""E2 lane area detectors (src/microsim/output/MSE2Collector.cpp) against the reference's own detector output.

The reference (`oracle/_ref/sumo -a e2.add.xml`) writes one <interval .../> per detector and period with 23 attributes (sampledSeconds,
nVehEntered, ..., jam lengths, halting durations, vehicle numbers); the engine's sumob200_e2_interval must give the same values --
compared as the strings the reference prints (two decimals), all of them, exactly: the engine adds the per-step notifications in
the reference's order (sorted by distance to the detector's end), so the sums are the same doubles.

Detectors: on every second lane one over the last 60 m before the junction (where queues form: jams, halting durations), on every
third lane one of 25 m from position 10 (vehicles depart on it), on every fifth lane one over the whole lane."""
import os
import re
import subprocess

import pytest

from conftest import CACHE, CASES, GOLDEN, REF, ensure_reference
from sumo_b200.flatfile import load_flat, load_trace

E2_CASES = ["grid4_tls", "grid5_dense", "c1_grid10", "grid4_tls_step05"]
PERIOD_S = 30
ATTRS = ["sampledSeconds", "nVehEntered", "nVehLeft", "nVehSeen", "meanSpeed", "meanTimeLoss", "meanOccupancy", "maxOccupancy",
         "meanMaxJamLengthInVehicles", "meanMaxJamLengthInMeters", "maxJamLengthInVehicles", "maxJamLengthInMeters", "jamLengthInVehiclesSum",
         "jamLengthInMetersSum", "meanHaltingDuration", "maxHaltingDuration", "haltingDurationSum", "meanIntervalHaltingDuration",
         "maxIntervalHaltingDuration", "intervalHaltingDurationSum", "startedHalts", "meanVehicleNumber", "maxVehicleNumber"]
INTEGER = {"nVehEntered", "nVehLeft", "nVehSeen", "maxJamLengthInVehicles", "jamLengthInVehiclesSum", "startedHalts", "maxVehicleNumber"}
# engine record field of each attribute
FIELD = {a: re.sub(r"(?<!^)([A-Z])", lambda m: "_" + m.group(1).lower(), a).replace("n_veh", "n_veh") for a in ATTRS}


def write_additional(net, path, out):
    lanes = re.findall(r'<lane id="([^":][^"]*)" index="0" speed="[^"]*" length="([^"]*)"', open(net).read())
    k = 0
    with open(path, "w") as f:
        f.write("<additional>\n")
        for i, (lane, length) in enumerate(lanes):
            length = float(length)
            defs = []
            if i % 2 == 0 and length > 70:
                defs.append('pos="%.2f" endPos="%.2f"' % (length - 60, length))
            if i % 3 == 0 and length > 40:
                defs.append('pos="10" length="25"')
            if i % 5 == 0:
                defs.append('pos="0" endPos="%.2f"' % length)
            for d in defs:
                f.write('  <laneAreaDetector id="a%04d" lane="%s" %s period="%d" file="%s"/>\n' % (k, lane, d, PERIOD_S, out))
                k += 1
        f.write("</additional>\n")
    return k


def e2_case(name):
    """(flat with the e2_* arrays, trace, {(begin string, detector id): {attribute: string}} of the reference's file)"""
    ensure_reference()
    os.makedirs(CACHE, exist_ok=True)
    base = name.replace("_step05", "")
    net, rou = os.path.join(GOLDEN, base + ".net.xml"), os.path.join(GOLDEN, base + ".rou.xml")
    add, out = os.path.join(CACHE, name + ".e2.add.xml"), os.path.join(CACHE, name + ".e2.out.xml")
    flat, trace = os.path.join(CACHE, name + ".e2.flat"), os.path.join(CACHE, name + ".e2.trace")
    exe = os.path.join(REF, "sumo_flatten")
    if not (os.path.exists(flat) and os.path.exists(out) and os.path.getmtime(flat) >= os.path.getmtime(exe)):
        assert write_additional(net, add, out) > 10
        common = ["-n", net, "-r", rou, "-a", add, "--no-step-log", "--no-warnings"] + CASES[name]
        subprocess.run([os.path.join(REF, "sumo")] + common, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        os.replace(out, out + ".ref")
        subprocess.run([exe, "--flat-out", flat, "--trace-out", trace] + common, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        os.replace(out + ".ref", out)
    ref = {}
    for m in re.finditer(r"<interval ([^>]*)/>", open(out).read()):
        a = dict(re.findall(r'(\w+)="([^"]*)"', m.group(1)))
        ref[(a["begin"], a["id"])] = a
    return flat, trace, ref


def detector_ids(flat):
    g = load_flat(flat)
    nb, names = g["e2_name_begin"], bytes(g["e2_names"]).decode()
    return [names[nb[i]:nb[i + 1]] for i in range(len(nb) - 1)]


def fmt(attr, value):
    return str(int(value)) if attr in INTEGER else "%.2f" % value


def compare(ref, got):
    """got: {(begin string, id): {attribute: number}}"""
    assert len(got) > 200 and set(got) <= set(ref)
    seen = jams = halts = 0
    for key, vals in got.items():
        want = ref[key]
        for a in ATTRS:
            assert fmt(a, vals[a]) == want[a], "%s: %s is %s, reference %s" % (key, a, fmt(a, vals[a]), want[a])
        seen += int(vals["nVehEntered"])
        jams += int(vals["jamLengthInVehiclesSum"])
        halts += int(vals["startedHalts"])
    assert seen > 500 and jams > 50 and halts > 50
    return seen


@pytest.mark.parametrize("case", E2_CASES)
def test_host_emulation_e2_matches_reference(case, host_emul, tmp_path):
    flat, trace, ref = e2_case(case)
    ids = detector_ids(flat)
    out = str(tmp_path / "e2.txt")
    n = len(load_trace(trace))
    subprocess.run([host_emul, flat, "--e2-out", out, "--e2-period", str(PERIOD_S * 1000), "--steps", str(n)], check=True, timeout=900, stderr=subprocess.DEVNULL)
    got = {}
    for line in open(out):
        f = line.split()
        vals = {a: float(x) for a, x in zip(ATTRS, f[3:])}
        got[("%.2f" % (int(f[0]) / 1000.), ids[int(f[2])])] = vals
    compare(ref, got)


@pytest.mark.gpu
@pytest.mark.parametrize("case", E2_CASES)
@pytest.mark.parametrize("reslot", [0, 7])
def test_engine_e2_matches_reference(case, reslot):
    from sumo_b200.engine import Engine
    flat, trace, ref = e2_case(case)
    ids = detector_ids(flat)
    n = len(load_trace(trace))
    got = {}
    with Engine(flat) as eng:
        eng.set_reslot_period(reslot)
        assert len(ids) == eng._lib.sumob200_e2_count(eng._h)
        for _ in range(n):
            r = eng.step(1)
            assert r.flags == 0
            if r.time_ms % (PERIOD_S * 1000) == 0:
                res = eng.e2_interval()
                for d, x in enumerate(res):
                    got[("%.2f" % (r.time_ms / 1000. - PERIOD_S), ids[d])] = {a: x[FIELD[a]] for a in ATTRS}
    compare(ref, got)
