import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = os.path.join(ROOT, "oracle", "_ref")
GOLDEN = os.path.join(ROOT, "tests", "golden")
CACHE = os.path.join(ROOT, "tests", "_cache")

# (fixture, extra reference options): every case is run through the UNMODIFIED reference (oracle/_ref/sumo_flatten) to get
# the flattened scenario and the ground-truth trace.  --default.speeddev 0 like tests/sumo/options.sumo of the reference.
CASES = {
    "grid3_small": ["--default.speeddev", "0"],
    "grid3_small_depart0": ["--default.speeddev", "0", "--default.departspeed", "0", "--default.departlane", "first"],
    "c1_grid10": ["--default.speeddev", "0", "--end", "3700"],
    "c1_grid10_depart0": ["--default.speeddev", "0", "--default.departspeed", "0", "--default.departlane", "first", "--end", "3700"],
    "grid4_tls": ["--default.speeddev", "0", "--end", "1200"],
    "grid5_dense": ["--default.speeddev", "0", "--end", "900"],
    # vehicles with individual speed factors (normc(1,0.1,0.2,2), sampled by the reference's loader)
    "grid4_tls_speeddev": ["--end", "1000"],
    # Krauss sigma = 0.5 (the vType default): the reference's per-vehicle dawdle draws are recorded (--draws-out) and replayed
    "grid4_tls_sigma": ["--default.speeddev", "0", "--end", "1200"],
    "grid5_dense_sigma": ["--default.speeddev", "0", "--end", "700"],
    "c1_grid10_sigma": ["--end", "2000"],
    # every vehicle carries MSDevice_Emissions: default class (HBEFA4/PC_petrol_Euro-4) and an explicit HBEFA3 class
    "grid4_tls_hbefa4": ["--default.speeddev", "0", "--end", "900", "--device.emissions.probability", "1"],
    "grid4_tls_hbefa3": ["--default.speeddev", "0", "--end", "900", "--device.emissions.probability", "1"],
    "grid5_dense_hbefa3": ["--default.speeddev", "0", "--end", "600", "--device.emissions.probability", "1"],
}
CASES.update({
    # explicit arrivalPos inside the last lane: vehicles arrive (and leave the lane queue) behind the front vehicle
    "grid4_tls_arrivalpos": ["--default.speeddev", "0", "--end", "1000"],
    # one district of bench.py's workload (10x10 TLS grid, 1800 resident vehicles, sigma 0.5, speedFactor deviation): same options as bench.py
    "bench_grid10_tls_draws": ["--default.departspeed", "0", "--default.departlane", "first", "--end", "700"],
    # carFollowModel IDM (MSCFModel_IDM, delta 4, 4 sub-steps)
    "grid3_small_idm": ["--default.speeddev", "0"],
    "grid5_dense_idm": ["--default.speeddev", "0", "--end", "900"],
    "c1_grid10_idm": ["--default.speeddev", "0", "--end", "2500"],
    # the case that exposed MSLink::blockedAtTime's LIVE foe speed under lastWasContRed (MSLink.cpp:1050) at t = 514 s: a left turner
    # waiting on the junction sees a foe whose lane the reference executes earlier in the step (executeMove pass 2, sb2_link.h)
    "grid4_tls_idm": ["--default.speeddev", "0", "--end", "1200"],
    # carFollowModel KraussOrig1 (MSCFModel_KraussOrig1: the SK safe speed): not collision free -- 20 resp. 58 rear-end collisions, each
    # resolved by teleporting the collider like the reference does (collision.action teleport)
    "grid4_tls_arrivalspeed": ["--default.speeddev", "0", "--end", "900"],
    "grid4_tls_orig1": ["--default.speeddev", "0", "--end", "900"],
    "grid5_dense_orig1": ["--default.speeddev", "0", "--end", "900"],
})
def _mixed_fleet(text):
    """three vTypes (default car, long slow truck, short quick car with impatience), assigned pseudo-randomly but reproducibly"""
    import random
    import re
    rnd = random.Random(7)
    text = text.replace('<vType id="car" sigma="0"/>', '''<vType id="car" sigma="0"/>
    <vType id="truck" sigma="0" length="12" minGap="3" accel="1.3" maxSpeed="25" tau="1.4" speedFactor="0.9"/>
    <vType id="fast" sigma="0" length="4.3" minGap="2" accel="3.5" maxSpeed="60" tau="1.2" impatience="0.3"/>''')
    return re.sub(r'type="car"', lambda m: 'type="%s"' % rnd.choice(["car", "car", "truck", "fast"]), text)


CASES.update({
    # other network shapes / junction types (netgenerate --rand, priority_stop, short traffic-light blocks with spill-back)
    "rand_net": ["--default.speeddev", "0", "--end", "900"],
    "grid4_stop": ["--default.speeddev", "0", "--end", "900"],
    "grid3_tls_short": ["--default.speeddev", "0", "--end", "900"],
    # other option combinations of the reference's command line
    "grid4_tls_step05": ["--default.speeddev", "0", "--step-length", "0.5", "--end", "600"],
    "grid5_dense_optmix": ["--default.speeddev", "0", "--default.departspeed", "desired", "--time-to-impatience", "30", "--end", "600"],
    # heterogeneous fleet: length, minGap, accel, maxSpeed, tau, speedFactor, impatience differ between vehicles
    # at t = 1086 s the advised speeds lead to a collision, which the reference resolves by teleporting the collider (collision.action
    # teleport, the default): the run goes on through it
    "grid4_tls_setspeed": ["--default.speeddev", "0", "--end", "1300"],
    "grid4_tls_setphase": ["--default.speeddev", "0", "--end", "1200"],
    "grid4_tls_mixed": ["--default.speeddev", "0", "--end", "1500"],
    "grid5_dense_mixed": ["--default.speeddev", "0", "--end", "900"],
    "c1_grid10_mixed": ["--default.speeddev", "0", "--end", "1500"],
})
# per-case edits of the fixture's route file: (old, new) or a function text -> text
ROU_EDITS = {
    "_mixed": _mixed_fleet,
    "_arrivalpos": ('<vehicle ', '<vehicle arrivalPos="60" '),
    "_idm": ('sigma="0"', 'sigma="0" carFollowModel="IDM"'),
    "_sigma": ('sigma="0"', 'sigma="0.5"'),
    "_orig1": ('sigma="0"', 'sigma="0" carFollowModel="KraussOrig1"'),
    # arrivalSpeed given: the vehicle brakes down to it towards its arrival position (MSVehicle.cpp:2752-2757; insertion: MSLane.cpp "arrival speed too low")
    "_arrivalspeed": ("<vehicle ", '<vehicle arrivalSpeed="3" '),
    "_hbefa3": ('sigma="0"', 'sigma="0" emissionClass="HBEFA3/PC_G_EU4"'),
}


def tls_phase_events(base, period_s=37, end_s=1100):
    """deterministic trafficlight.setPhase events for the case <base>_setphase: every period_s seconds one program is set to some
    phase (sometimes the one it is in: only the switch is rescheduled).  [(time ms, tls id, phase)]"""
    import re
    text = open(os.path.join(GOLDEN, base + ".net.xml")).read()
    logics = [(m.group(1), len(re.findall(r"<phase ", m.group(2)))) for m in re.finditer(r'<tlLogic id="([^"]+)"[^>]*>(.*?)</tlLogic>', text, re.S)]
    logics = [l for l in logics if l[1] > 1]
    ev = []
    for k, t in enumerate(range(period_s, end_s, period_s)):
        tid, n = logics[(k * 7) % len(logics)]
        ev.append((t * 1000, tid, (k * 3) % n))
        if k % 4 == 0:   # two programs in the same step
            tid2, n2 = logics[(k * 7 + 3) % len(logics)]
            ev.append((t * 1000, tid2, (k + 1) % n2))
    return ev


def veh_speed_events(base, end_s=1290):
    """deterministic vehicle.setSpeed events for the case <base>_setspeed: every 5 s two vehicles get a speed advice (0 = stop,
    slow, medium, above the lane limit, or -1 = release), slow ones are released 30 s later.  Both sides skip an event whose
    vehicle is not on the road at that time.  [(time ms, vehicle id, speed)]"""
    import re
    ids = re.findall(r'<vehicle id="([^"]+)"', open(os.path.join(GOLDEN, base + ".rou.xml")).read())
    speeds = [0., 3., 7.5, 20., -1., 11.]
    ev = []
    for k, t in enumerate(range(50, end_s, 5)):
        for j in range(2):
            # vehicles depart in id order over the run: aim at those that left a little while ago
            i = max(0, min(int(len(ids) * t / 900.), len(ids) - 1) - 20 - ((k * 13 + j * 29) % 60))
            sp = speeds[(k * 2 + j) % len(speeds)]
            ev.append((t * 1000, ids[i], sp))
            if 0 <= sp < 5:
                ev.append(((t + 30) * 1000, ids[i], -1.))
    return sorted(ev, key=lambda e: e[0])


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def ensure_reference():
    """oracle/_ref/sumo_flatten: prebuilt on the GPU box, built from /root/reference in the build container"""
    exe = os.path.join(REF, "sumo_flatten")
    if not os.path.exists(exe):
        if not os.path.isdir(os.environ.get("SUMO_REFERENCE", "/root/reference")):
            pytest.skip("oracle/_ref/sumo_flatten missing and no reference sources to build it from")
        subprocess.check_call([os.path.join(ROOT, "oracle", "build_ref.sh"), "flatten"])
    return exe


def draws_path(name):
    return os.path.join(CACHE, name + ".draws")


def emissions_path(name):
    return os.path.join(CACHE, name + ".emis")


def load_emissions(path):
    """per step (t_ms, vehicle numerical ids, [n,7] accumulated CO2 CO HC fuel NOx PMx electricity) of the on-road vehicles:
    MSDevice_Emissions::myEmissions of the reference (sumo_flatten --emissions-out) or of the host emulation"""
    import numpy as np
    raw = np.fromfile(path, dtype=np.int64)
    out, off = [], 0
    while off < raw.size:
        t, n = int(raw[off]), int(raw[off + 1])
        rec = raw[off + 2:off + 2 + 8 * n].reshape(n, 8)
        out.append((t, rec[:, 0].copy(), rec[:, 1:].copy().view(np.float64)))
        off += 2 + 8 * n
    return out


def load_draws(path):
    """per step (t_ms, vehicle numerical ids, draws, prediction verified) written by sumo_flatten --draws-out"""
    import numpy as np
    raw = np.fromfile(path, dtype=np.int64)
    out, off = [], 0
    while off < raw.size:
        t, n, ok = int(raw[off]), int(raw[off + 1]), int(raw[off + 2])
        rec = raw[off + 3:off + 3 + 2 * n].reshape(n, 2)
        out.append((t, rec[:, 0].copy(), rec[:, 1].copy().view(np.float64), ok))
        off += 3 + 2 * n
    return out


def reference_case(name):
    """returns (flat path, trace path) for a case, generating them with the reference binary on first use"""
    exe = ensure_reference()
    os.makedirs(CACHE, exist_ok=True)
    flat = os.path.join(CACHE, name + ".flat")
    trace = os.path.join(CACHE, name + ".trace")
    if not (os.path.exists(flat) and os.path.exists(trace) and os.path.getmtime(flat) >= os.path.getmtime(exe)):
        base = name
        for suffix in ("_depart0", "_speeddev", "_teleport", "_sigma", "_hbefa4", "_hbefa3", "_idm", "_orig1", "_arrivalspeed", "_draws", "_arrivalpos", "_mixed", "_step05", "_optmix", "_setphase", "_setspeed"):
            base = base.replace(suffix, "")
        rou = os.path.join(GOLDEN, base + ".rou.xml")
        extra = []
        for suffix, edit in ROU_EDITS.items():
            if name.endswith(suffix):
                text = open(rou).read()
                if callable(edit):
                    new_text = edit(text)
                else:
                    assert edit[0] in text
                    new_text = text.replace(edit[0], edit[1])
                assert new_text != text
                rou = os.path.join(CACHE, name + ".rou.xml")
                with open(rou, "w") as f:
                    f.write(new_text)
        if name.endswith("_setspeed"):
            extra = ["--veh-set-speed", ";".join("%d:%s:%.17g" % e for e in veh_speed_events(base))]
        if name.endswith("_setphase"):
            extra = ["--tls-set-phase", ";".join("%d:%s:%d" % e for e in tls_phase_events(base))]
        if name.endswith(("_sigma", "_draws")):
            extra = ["--draws-out", draws_path(name)]
        if "_hbefa" in name:
            extra = ["--emissions-out", emissions_path(name)]
        cmd = [exe, "--flat-out", flat, "--trace-out", trace + ".tmp", "-n", os.path.join(GOLDEN, base + ".net.xml"),
               "-r", rou, "--no-step-log", "--no-warnings"] + extra + CASES[name]
        subprocess.run(cmd, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        os.replace(trace + ".tmp", trace)
    return flat, trace


def reference_fcd(name):
    """{(time string, vehicle id): (x, y, angle, speed, pos, lane id)} from the reference's own --fcd-output (2 decimals)"""
    import re
    exe = os.path.join(REF, "sumo")
    ensure_reference()
    os.makedirs(CACHE, exist_ok=True)
    out = os.path.join(CACHE, name + ".fcd.xml")
    if not os.path.exists(out):
        cmd = [exe, "-n", os.path.join(GOLDEN, name + ".net.xml"), "-r", os.path.join(GOLDEN, name + ".rou.xml"), "--no-step-log",
               "--no-warnings", "--fcd-output", out + ".tmp"] + CASES[name]
        subprocess.run(cmd, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        os.replace(out + ".tmp", out)
    ref, t = {}, None
    veh = re.compile(r'<vehicle id="([^"]+)" x="([^"]+)" y="([^"]+)" angle="([^"]+)" type="[^"]*" speed="([^"]+)" pos="([^"]+)" lane="([^"]+)"')
    for line in open(out):
        m = re.search(r'<timestep time="([0-9.]+)"', line)
        if m:
            t = m.group(1)
            continue
        m = veh.search(line)
        if m:
            ref[(t, m.group(1))] = m.groups()[1:]
    return ref


def reference_fcd_body(name):
    """the <timestep> ... </timestep> lines of the reference's --fcd-output file, as bytes"""
    reference_fcd(name)
    out = os.path.join(CACHE, name + ".fcd.xml")
    return b"".join(l for l in open(out, "rb") if b"timestep" in l or b"<vehicle" in l)


def reference_tripinfo(name):
    """{vehicle id: dict of the attributes of its <tripinfo>} from the reference's own --tripinfo-output"""
    import re
    exe = os.path.join(REF, "sumo")
    ensure_reference()
    os.makedirs(CACHE, exist_ok=True)
    out = os.path.join(CACHE, name + ".tripinfo.xml")
    base, rou = name, os.path.join(GOLDEN, name + ".rou.xml")
    if not os.path.exists(rou):
        # a case with an edited route file (ROU_EDITS): reference_case(name) wrote it into the cache
        reference_case(name)
        rou = os.path.join(CACHE, name + ".rou.xml")
        for suffix in ROU_EDITS:
            base = base.replace(suffix, "")
    if not os.path.exists(out):
        cmd = [exe, "-n", os.path.join(GOLDEN, base + ".net.xml"), "-r", rou, "--no-step-log",
               "--no-warnings", "--tripinfo-output", out + ".tmp"] + CASES[name]
        subprocess.run(cmd, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        os.replace(out + ".tmp", out)
    ref = {}
    for line in open(out):
        if "<tripinfo " in line:
            a = dict(re.findall(r'(\w+)="([^"]*)"', line))
            ref[a["id"]] = a
    return ref


def reference_summary(name):
    """[dict of the attributes of every <step> of the reference's own --summary-output]; cases whose suffix only adds options"""
    import re
    exe = os.path.join(REF, "sumo")
    ensure_reference()
    os.makedirs(CACHE, exist_ok=True)
    out = os.path.join(CACHE, name + ".summary.xml")
    base = name
    for suffix in ("_depart0", "_step05", "_optmix"):
        base = base.replace(suffix, "")
    if not os.path.exists(out):
        cmd = [exe, "-n", os.path.join(GOLDEN, base + ".net.xml"), "-r", os.path.join(GOLDEN, base + ".rou.xml"), "--no-step-log",
               "--no-warnings", "--summary-output", out + ".tmp"] + CASES[name]
        subprocess.run(cmd, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        os.replace(out + ".tmp", out)
    return [dict(re.findall(r'(\w+)="([^"]*)"', line)) for line in open(out) if "<step " in line]


def summary_strings(rec):
    """a summary record (engine SUMMARY row or the same fields from the host emulation) formatted like MSNet::writeSummaryOutput"""
    return {"time": "%.2f" % (int(rec["time_ms"]) / 1000.), "inserted": str(int(rec["inserted"])), "running": str(int(rec["running"])),
            "waiting": str(int(rec["waiting"])), "ended": str(int(rec["ended"])), "arrived": str(int(rec["arrived"])),
            "collisions": str(int(rec["collisions"])), "teleports": str(int(rec["teleports"])), "halting": str(int(rec["halting"])),
            "stopped": str(int(rec["stopped"])), "meanWaitingTime": "%.2f" % rec["mean_waiting_time"], "meanTravelTime": "%.2f" % rec["mean_travel_time"],
            "meanSpeed": "%.2f" % rec["mean_speed"], "meanSpeedRelative": "%.2f" % rec["mean_speed_relative"]}


def tripinfo_strings(rec, lane_names):
    """a trip record (engine TRIPINFO row or the same fields from the host emulation) formatted like MSDevice_Tripinfo::generateOutput
    with the default precision 2 (time2string rounds the ms value half up)"""
    def t(ms):
        return "%.2f" % (((int(ms) + 5) // 10) / 100.)
    return {"depart": t(rec["depart_ms"]), "departLane": lane_names[int(rec["depart_lane"])], "departPos": "%.2f" % rec["depart_pos"],
            "departSpeed": "%.2f" % rec["depart_speed"], "departDelay": t(rec["depart_delay_ms"]), "arrival": t(rec["arrival_ms"]),
            "arrivalLane": lane_names[int(rec["arrival_lane"])], "arrivalPos": "%.2f" % rec["arrival_pos"], "arrivalSpeed": "%.2f" % rec["arrival_speed"],
            "duration": t(rec["duration_ms"]), "routeLength": "%.2f" % rec["route_length"], "waitingTime": t(rec["waiting_ms"]),
            "waitingCount": str(int(rec["waiting_count"])), "timeLoss": t(rec["time_loss_ms"]), "speedFactor": "%.2f" % rec["speed_factor"]}


@pytest.fixture(scope="session")
def host_emul():
    """tests/host_emul: the device functions compiled for the host (test infrastructure, see its header)"""
    os.makedirs(CACHE, exist_ok=True)
    exe = os.path.join(CACHE, "host_emul")
    src = os.path.join(ROOT, "tests", "host_emul", "emul_main.cpp")
    deps = [src] + [os.path.join(ROOT, "sumo_b200", "csrc", f) for f in os.listdir(os.path.join(ROOT, "sumo_b200", "csrc"))]
    if not os.path.exists(exe) or any(os.path.getmtime(d) > os.path.getmtime(exe) for d in deps):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-pthread", "-o", exe, src])
    return exe
