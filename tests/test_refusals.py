"""Scenarios with features the step engine does not model are REFUSED when the scenario is flattened (integration/sumo_flatten.cpp)
or when the engine is created (sumo_b200/csrc/sb2_build.h) -- with a message, never simplified silently."""
import os
import subprocess

import pytest

from conftest import GOLDEN, ensure_reference

NET = os.path.join(GOLDEN, "grid3_small.net.xml")


def first_route():
    import re
    text = open(os.path.join(GOLDEN, "grid3_small.rou.xml")).read()
    m = re.search(r'<route edges="([^"]+)"', text)
    return m.group(1)


ROUTE_FILES = {
    "stop": ('<routes><vType id="car" sigma="0"/><vehicle id="a" type="car" depart="0"><route edges="%(edges)s"/>'
             '<stop lane="%(first)s_0" endPos="20" duration="10"/></vehicle></routes>', "<stop>s are not supported"),
    "person": ('<routes><vType id="car" sigma="0"/><vehicle id="a" type="car" depart="0"><route edges="%(edges)s"/></vehicle>'
               '<person id="p" depart="0"><walk edges="%(first)s"/></person></routes>', "persons / containers are not supported"),
    "arrivalLane": ('<routes><vType id="car" sigma="0"/><vehicle id="a" type="car" depart="0" arrivalLane="random"><route edges="%(edges)s"/></vehicle></routes>',
                    "arrivalLane is not supported"),
    "actionStepLength": ('<routes><vType id="car" sigma="0" actionStepLength="2"/><vehicle id="a" type="car" depart="0"><route edges="%(edges)s"/></vehicle></routes>',
                         "actionStepLength"),
    "jm": ('<routes><vType id="car" sigma="0" jmIgnoreFoeProb="0.5"/><vehicle id="a" type="car" depart="0"><route edges="%(edges)s"/></vehicle></routes>',
           "junction model parameters"),
}


@pytest.mark.parametrize("feature", sorted(ROUTE_FILES))
def test_flatten_refuses(feature, tmp_path):
    exe = ensure_reference()
    edges = first_route()
    text, message = ROUTE_FILES[feature]
    rou = str(tmp_path / "r.rou.xml")
    open(rou, "w").write(text % {"edges": edges, "first": edges.split()[0]})
    r = subprocess.run([exe, "--flat-out", str(tmp_path / "x.flat"), "-n", NET, "-r", rou, "--no-step-log"], capture_output=True, text=True, timeout=120)
    assert r.returncode != 0, "the scenario was accepted"
    assert message in (r.stdout + r.stderr), (r.stdout + r.stderr)[-800:]
    assert not os.path.exists(str(tmp_path / "x.flat"))


FLOW = '<routes><vType id="car" sigma="0"/><route id="r" edges="%s"/><flow id="f" type="car" route="r" begin="0" end="100" %s/></routes>'


@pytest.mark.parametrize("spec", ['period="10"', 'number="7"', 'probability="0.2"', 'period="exp(0.1)"'])
def test_flows_are_expanded_by_the_harness_as_the_reference_expands_them(spec, tmp_path, host_emul):
    """<flow>: the parity harness lets the reference's own MSInsertionControl::determineCandidates build the vehicles before the run
    (integration/sumo_flatten.cpp expandFlows).  The expanded run must be the untouched reference's run (trace without --flat-out),
    and the engine must reproduce it."""
    exe = ensure_reference()
    rou = str(tmp_path / "f.rou.xml")
    open(rou, "w").write(FLOW % (first_route(), spec))
    common = ["-n", NET, "-r", rou, "--no-step-log", "--no-warnings", "--seed", "7"]
    subprocess.run([exe, "--trace-out", str(tmp_path / "n.trace")] + common, check=True, capture_output=True, timeout=120)
    subprocess.run([exe, "--flat-out", str(tmp_path / "x.flat"), "--trace-out", str(tmp_path / "x.trace")] + common, check=True, capture_output=True, timeout=120)
    natural = open(str(tmp_path / "n.trace"), "rb").read()
    assert len(natural) > 5000 and natural == open(str(tmp_path / "x.trace"), "rb").read()
    from sumo_b200.flatfile import load_trace
    steps = len(load_trace(str(tmp_path / "x.trace")))
    subprocess.run([host_emul, str(tmp_path / "x.flat"), "--trace-out", str(tmp_path / "e.trace"), "--steps", str(steps)], check=True, capture_output=True, timeout=120)
    assert natural == open(str(tmp_path / "e.trace"), "rb").read()


def test_hook_refuses_flows(tmp_path):
    """the drop-in binary loads everything at once and does not expand flows: refused with a message (no GPU needed: the scenario is
    refused before the engine is created)"""
    ensure_reference()
    from conftest import REF
    exe = os.path.join(REF, "sumo_b200")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/sumo_b200 is not built")
    rou = str(tmp_path / "f.rou.xml")
    open(rou, "w").write(FLOW % (first_route(), 'period="10"'))
    r = subprocess.run([exe, "--b200.device", "0", "-n", NET, "-r", rou, "--no-step-log"], capture_output=True, text=True, timeout=120)
    assert "<flow> definitions are not supported" in (r.stdout + r.stderr)


def test_rerouting_device_is_refused(tmp_path):
    exe = ensure_reference()
    rou = os.path.join(GOLDEN, "grid3_small.rou.xml")
    r = subprocess.run([exe, "--flat-out", str(tmp_path / "x.flat"), "-n", NET, "-r", rou, "--no-step-log", "--device.rerouting.probability", "1"],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode != 0 and "device 'rerouting' is not supported" in (r.stdout + r.stderr)


def test_plain_scenario_is_accepted(tmp_path):
    exe = ensure_reference()
    r = subprocess.run([exe, "--flat-out", str(tmp_path / "x.flat"), "-n", NET, "-r", os.path.join(GOLDEN, "grid3_small.rou.xml"), "--no-step-log"],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and os.path.exists(str(tmp_path / "x.flat"))


def test_depart_position_inside_the_first_vehicle_length_is_refused_by_the_engine(tmp_path, host_emul):
    """departPos < vehicle length: the vehicle would lap into the predecessor lanes at insertion (MSVehicle::computeFurtherLanes) --
    refused when the engine's model is built (sb2_build.h; the host emulation runs the same builder); 6 m (> 5 m) is accepted"""
    exe = ensure_reference()
    edges = first_route()
    for pos, ok in (("2", False), ("6", True)):
        rou = str(tmp_path / ("r%s.rou.xml" % pos))
        open(rou, "w").write('<routes><vType id="car" sigma="0"/><vehicle id="a" type="car" depart="0" departPos="%s"><route edges="%s"/></vehicle></routes>' % (pos, edges))
        flat = str(tmp_path / ("x%s.flat" % pos))
        subprocess.run([exe, "--flat-out", flat, "-n", NET, "-r", rou, "--no-step-log", "--end", "20"], check=True, capture_output=True, timeout=120)
        r = subprocess.run([host_emul, flat, "--steps", "5"], capture_output=True, text=True, timeout=120)
        if ok:
            assert r.returncode == 0, r.stderr[-500:]
        else:
            assert r.returncode != 0 and "departPos closer to the lane's start" in r.stderr
