"""dev helper: profiles/r02_texttest_sweep.tsv from the raw output(s) of tools_dev/texttest_sweep.py (later files override earlier
ones test by test): header with the totals and the histogram of refusal reasons, then the tests that ran (SAME / DIFF / FLAGGED /
FLOWDIFF lines in full; the refused, failed-to-load and empty ones are only counted).
Usage: sweep_report.py OUT.tsv RAW.tsv [RAW2.tsv ...]"""
import re
import sys

rows = {}
for path in sys.argv[2:]:
    for line in open(path):
        f = line.rstrip("\n").split("\t")
        if len(f) >= 2 and not line.startswith("#"):
            rows[f[1]] = (f[0], f[2] if len(f) > 2 else "")
counts, reasons = {}, {}
for rel, (status, info) in rows.items():
    counts[status] = counts.get(status, 0) + 1
    if status == "REFUSED":
        msg = re.sub(r"'[^']*'", "'..'", info[info.find("sumob200"):])
        reasons[msg] = reasons.get(msg, 0) + 1
vehsteps = sum(int(re.search(r"(\d+) vehicle-steps", info).group(1)) for s, info in rows.values() if s == "SAME")
with open(sys.argv[1], "w") as out:
    out.write("# tools_dev/texttest_sweep.py over /root/reference/tests/sumo/{action_step_length,basic,bugs,cf_model,devices,errors,extended,junction_model,"
              "lc_model,meta,opposite_direction_driving,output,pedestrian_model,rail,spec,sublane_model,tls}\n")
    out.write("# every test directory: the reference's loader flattens the scenario (oracle/_ref/sumo_flatten; <flow>s expanded up front by the reference's own\n"
              "# MSInsertionControl::determineCandidates, checked against the unexpanded run), tests/host_emul replays it; SAME = lane, position and speed of every\n"
              "# vehicle after every step bit-identical to the reference's own run (dawdle draws replayed); REFUSED = the scenario uses something the engine refuses\n"
              "# (message); FLAGGED = a fixed capacity of the engine was exceeded during the run (SUMOB200_F_*_OVERFLOW raised: the run is void by contract);\n"
              "# LOADFAIL = the reference itself rejects the options (tests of error messages, missing inputs of this build); EMPTY = no vehicle on the road;\n"
              "# SKIP = TraCI / config-file tests\n")
    out.write("# totals: " + ", ".join("%s %d" % (k, counts[k]) for k in sorted(counts)) + "; SAME covers %d vehicle-steps\n" % vehsteps)
    out.write("# refusal reasons:\n")
    for msg, n in sorted(reasons.items(), key=lambda kv: -kv[1]):
        out.write("# %5d  %s\n" % (n, msg))
    for rel in sorted(rows):
        status, info = rows[rel]
        if status not in ("REFUSED", "LOADFAIL", "EMPTY", "SKIP"):
            out.write("%s\t%s\t%s\n" % (status, rel, info))
print(counts, vehsteps)
