"""dev helper: the engine against the reference's own texttest scenarios, in bulk.  For every test directory below the given
sub-trees of /root/reference/tests/sumo (a directory with an options.sumo and no sub-directories that have one): texttest's option
inheritance, the reference's loader flattens the scenario and records the reference's trajectories (oracle/_ref/sumo_flatten
--trace-out, --draws-out), tests/host_emul runs the same steps with the draws replayed, the two traces are compared byte by byte.
The sweep compares trajectories only: sumo_flatten --trajectories-only leaves observers the engine has no counterpart for (SSM / bluetooth /
battery devices, E3 and other detectors, emission / noise mean data) out of the flat file instead of refusing the scenario.
Outcomes: SAME, DIFF (first differing step), REFUSED (the message), LOADFAIL (the reference itself rejects the options, e.g. tests of
error messages), EMPTY (no vehicle ever on the road), FLAGGED (the engine raised one of its capacity flags during the run).
FLOWDIFF would mean the harness's up-front <flow> expansion changed the reference's own trajectories (never seen).
Usage: texttest_sweep.py OUT.tsv SUBTREE|@LISTFILE [...]     e.g.  basic cf_model junction_model tls"""
import os
import shlex
import shutil
import subprocess
import sys
import tempfile
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sumo_b200.flatfile import load_trace  # noqa: E402

TT_ROOT = "/root/reference/tests/sumo"
REF = os.path.join(ROOT, "oracle", "_ref")
EMUL = os.path.join(ROOT, "tests", "_cache", "host_emul")


def stage(rel, tmp):
    parts = rel.split("/")
    dirs = [TT_ROOT] + [os.path.join(TT_ROOT, *parts[:i + 1]) for i in range(len(parts))]
    options = []
    for d in dirs:
        f = os.path.join(d, "options.sumo")
        if os.path.exists(f):
            options += shlex.split(open(f).read())
        for fn in os.listdir(d):
            p = os.path.join(d, fn)
            if os.path.isfile(p) and not fn.endswith(".sumo") and not fn.startswith("testsuite") and ".sumo." not in fn and os.path.getsize(p) < 50e6:
                shutil.copy(p, os.path.join(tmp, fn))
    return options


def has_flows(tmp):
    for fn in os.listdir(tmp):
        if fn.endswith(".xml"):
            with open(os.path.join(tmp, fn), "rb") as f:
                if b"<flow" in f.read():
                    return True
    return False


def run_one(rel):
    os.makedirs("/tmp/w", exist_ok=True)
    tmp = tempfile.mkdtemp(prefix="tt_", dir="/tmp/w")
    try:
        options = stage(rel, tmp)
        if any(o.startswith("--remote-port") or o in ("-c", "--configuration-file") for o in options):
            return rel, "SKIP", "traci / config file"
        try:
            r = subprocess.run([os.path.join(REF, "sumo_flatten"), "--trajectories-only", "--flat-out", "x.flat", "--trace-out", "x.trace", "--draws-out", "x.draws"] + options,
                               cwd=tmp, capture_output=True, text=True, timeout=120)
        except subprocess.TimeoutExpired:
            return rel, "TIMEOUT", "reference > 120 s"
        if r.returncode != 0:
            err = (r.stderr + r.stdout).strip().splitlines()
            msg = next((l for l in err if "sumob200" in l), err[-1] if err else "")
            return rel, ("REFUSED" if "sumob200" in msg else "LOADFAIL"), msg[:160]
        if not os.path.exists(os.path.join(tmp, "x.trace")):
            return rel, "LOADFAIL", "no trace"
        ref = load_trace(os.path.join(tmp, "x.trace"))
        if has_flows(tmp):
            # the harness expanded <flow>s before the run (integration/sumo_flatten.cpp expandFlows): the untouched reference,
            # which expands them step by step, must give the same trace
            n = subprocess.run([os.path.join(REF, "sumo_flatten"), "--trace-out", "n.trace"] + options, cwd=tmp, capture_output=True, text=True, timeout=120)
            if n.returncode != 0 or open(os.path.join(tmp, "x.trace"), "rb").read() != open(os.path.join(tmp, "n.trace"), "rb").read():
                return rel, "FLOWDIFF", "expanded flows change the reference's own run"
        if sum(len(s[1]) for s in ref) == 0:
            return rel, "EMPTY", ""
        cmd = [EMUL, "x.flat", "--trace-out", "e.trace", "--draws", "x.draws", "--steps", str(len(ref))]
        note = ""
        try:
            e = subprocess.run(cmd, cwd=tmp, capture_output=True, text=True, timeout=300)
            if e.returncode == 0 and "err=" in e.stderr and int(e.stderr.strip().splitlines()[-1].rsplit("err=", 1)[-1], 16) & 1:
                # more than 8 links inside the look-ahead (very short lanes): the scenario asks for more drive items per vehicle
                e = subprocess.run(cmd + ["--drive-items", "32"], cwd=tmp, capture_output=True, text=True, timeout=300)
                note = " (drive_item_capacity 32)"
        except subprocess.TimeoutExpired:
            return rel, "TIMEOUT", "emulation > 300 s"
        if e.returncode != 0:
            msg = (e.stderr.strip().splitlines() or [""])[-1]
            return rel, ("REFUSED" if "sumob200" in msg else "EMULFAIL"), msg[:160]
        if open(os.path.join(tmp, "x.trace"), "rb").read() == open(os.path.join(tmp, "e.trace"), "rb").read():
            return rel, "SAME", "%d steps, %d vehicle-steps%s" % (len(ref), sum(len(s[1]) for s in ref), note)
        flags = int(e.stderr.strip().splitlines()[-1].rsplit("err=", 1)[-1], 16) if "err=" in e.stderr else 0
        if flags & (1 | 2 | 4 | 512 | 8192):
            # a fixed capacity of the engine was exceeded (drive items, further lanes, partial occupators, link leaders, detector slots):
            # the run is void by contract (SUMOB200_F_*_OVERFLOW), a refusal at run time
            return rel, "FLAGGED", "capacity flag 0x%x" % flags
        mine = load_trace(os.path.join(tmp, "e.trace"))
        for i, (a, b) in enumerate(zip(ref, mine)):
            if len(a[1]) != len(b[1]) or (a[1] != b[1]).any() or (a[2] != b[2]).any() or (a[3] != b[3]).any() or (a[4] != b[4]).any():
                return rel, "DIFF", "step %d of %d; %s" % (i, len(ref), (e.stderr.strip().splitlines() or [""])[-1][-80:])
        return rel, "DIFF", "length %d vs %d" % (len(ref), len(mine))
    except Exception as ex:  # noqa: BLE001
        return rel, "ERROR", repr(ex)[:160]
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def main():
    out = sys.argv[1]
    tests = []
    for sub in sys.argv[2:]:
        if sub.startswith("@"):      # a file with one test directory per line
            tests += [l.strip() for l in open(sub[1:]) if l.strip()]
            continue
        for d, subdirs, files in os.walk(os.path.join(TT_ROOT, sub)):
            if "options.sumo" in files and not any(os.path.exists(os.path.join(d, s, "options.sumo")) for s in subdirs):
                tests.append(os.path.relpath(d, TT_ROOT))
    tests.sort()
    print(len(tests), "tests", flush=True)
    counts = {}
    with ThreadPoolExecutor(max_workers=8) as pool, open(out, "w") as f:
        for rel, status, info in pool.map(run_one, tests):
            counts[status] = counts.get(status, 0) + 1
            f.write("%s\t%s\t%s\n" % (status, rel, info))
            f.flush()
    print(counts)


if __name__ == "__main__":
    main()
