#!/usr/bin/env python3
"""Generates the 5 x 5 template network of sumo_b200/gridgen.py (run in the build container, needs oracle/_ref):

  grid5_tls800.flat   netgenerate --grid --grid.number=5 --grid.length=800 --default-junction-type traffic_light (the UNMODIFIED
                      reference netgenerate), loaded and flattened by the reference's own loader (oracle/_ref/sumo_flatten) with one
                      dummy vehicle of the default vType (gridgen.make_demand replaces the demand).

bench.py instantiates it as the 100 x 100 grid of BASELINE.json configs[1].
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "..", "..", "oracle", "_ref")


def make(name, number, length, junction_type, extra=()):
    net = os.path.join("/tmp", name + ".net.xml")
    rou = os.path.join("/tmp", name + ".rou.xml")
    subprocess.check_call([os.path.join(REF, "netgenerate"), "--grid", "--grid.number=%d" % number, "--grid.length=%d" % length,
                           "--default-junction-type", junction_type, "-o", net, "--no-warnings"] + list(extra),
                          stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    with open(rou, "w") as f:
        f.write('<routes><vType id="car" sigma="0.5"/><vehicle id="v0" type="car" depart="0"><route edges="A0A1 A1A2"/></vehicle></routes>\n')
    flat = os.path.join(HERE, name + ".flat")
    subprocess.check_call([os.path.join(REF, "sumo_flatten"), "--flat-out", flat, "-n", net, "-r", rou, "--no-step-log", "--no-warnings",
                           "--default.departspeed", "0", "--default.departlane", "first", "--device.emissions.probability", "1"],
                          stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    print(flat, os.path.getsize(flat))


if __name__ == "__main__":
    make("grid5_tls800", 5, 800, "traffic_light")
    if "--all" in sys.argv:
        make("grid5_prio100", 5, 100, "priority")
