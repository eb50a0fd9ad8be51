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
    "flow": ('<routes><vType id="car" sigma="0"/><route id="r" edges="%(edges)s"/><flow id="f" type="car" route="r" begin="0" end="100" period="10"/></routes>',
             "<flow> definitions are not supported"),
    "person": ('<routes><vType id="car" sigma="0"/><vehicle id="a" type="car" depart="0"><route edges="%(edges)s"/></vehicle>'
               '<person id="p" depart="0"><walk edges="%(first)s"/></person></routes>', "persons / containers are not supported"),
    "arrivalLane": ('<routes><vType id="car" sigma="0"/><vehicle id="a" type="car" depart="0" arrivalLane="0"><route edges="%(edges)s"/></vehicle></routes>',
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
