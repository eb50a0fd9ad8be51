"""Pins the oracle build to the stock build: oracle/_ref/sumo (the reference's sources compiled here against the expat-backed
Xerces shim, oracle/build_ref.sh) must reproduce the reference's OWN texttest goldens (tests/sumo/**/fcd.sumo, aggregated60.sumo:
files the reference's maintainers committed from a stock cmake + Xerces build) on the reference's own inputs and options.

texttest semantics restated: the options of a test are the concatenation of the options.sumo files from tests/sumo down to the
test directory; input files are looked up in the test directory and its ancestors (the nearest wins).

Needs /root/reference (the build container); skipped where it is absent (the GPU box only gets the prebuilt binaries)."""
import os
import re
import shlex
import shutil
import subprocess

import pytest

from conftest import REF

REFERENCE = os.environ.get("SUMO_REFERENCE", "/root/reference")
TT_ROOT = os.path.join(REFERENCE, "tests", "sumo")

# (test directory below tests/sumo, file the run writes, committed golden)
TEXTTESTS = [
    ("output/fcd-output/begin", "fcd.xml", "fcd.sumo"),
    ("output/fcd-output/period", "fcd.xml", "fcd.sumo"),
    ("output/fcd-output/subsecond", "fcd.xml", "fcd.sumo"),
    ("output/fcd-output/angle", "fcd.xml", "fcd.sumo"),
    ("output/fcd-output/acceleration", "fcd.xml", "fcd.sumo"),
    ("output/fcd-output/short_lane", "fcd.xml", "fcd.sumo"),
    # Krauss / IDM car following on a circle, 3000 s, lane-based mean data every 60 s (speed, density, waiting time ... at 2 decimals)
    ("cf_model/drive_in_circles/krauss", "aggregated_60.xml", "aggregated60.sumo"),
    ("cf_model/drive_in_circles/idm", "aggregated_60.xml", "aggregated60.sumo"),
    # two lanes: MSLaneChanger + LC2013 of the oracle build
    ("cf_model/drive_in_circles/circle_two_lanes", "aggregated_60.xml", "aggregated60.sumo"),
]


def xml_body(path):
    text = re.sub(r"<!--.*?-->", "", open(path).read(), flags=re.S)   # the header comment carries the build version and date
    return [l.strip() for l in text.splitlines() if l.strip() and not l.strip().startswith("<?xml")]


@pytest.mark.parametrize("rel,produced,golden", TEXTTESTS)
def test_oracle_build_reproduces_reference_golden(rel, produced, golden, tmp_path):
    if not os.path.isdir(TT_ROOT):
        pytest.skip("reference sources (tests/sumo) are not available on this machine")
    exe = os.path.join(REF, "sumo")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/sumo is not built")
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
    r = subprocess.run([exe] + options, cwd=str(tmp_path), capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    got = xml_body(str(tmp_path / produced))
    want = xml_body(os.path.join(TT_ROOT, rel, golden))
    assert len(want) > 5
    assert got == want
