#!/usr/bin/env python3
"""Generates the base tile of the throughput benchmark (run in the build container, needs oracle/_ref):

  bench_grid10_tls.net.xml   netgenerate --grid --grid.number=10 --grid.length=200 --default-junction-type traffic_light
                             (the UNMODIFIED reference netgenerate; static TLS programs as netgenerate builds them)
  bench_grid10_tls.rou.xml   N vehicles, Krauss, random-walk routes of ROUTE_EDGES edges without U-turns (python random,
                             seeded), depart uniform in [0, DEPART_END) s  -> vehicles stay resident for the measured window
  bench_grid10_tls.flat      the scenario flattened by the reference's own loader (oracle/_ref/sumo_flatten)

bench.py tiles the .flat into as many disjoint districts as the requested resident vehicle count needs (sumo_b200/tiler.py).
"""
import os
import random
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_fixtures import read_graph  # noqa: E402

REF = os.path.join(HERE, "..", "..", "oracle", "_ref")
NAME = "bench_grid10_tls"
N_VEH = 1800          # 72 lane-km -> 25 veh/lane-km
ROUTE_EDGES = 70
DEPART_END = 300.0
SEED = 42


def main():
    net = os.path.join(HERE, NAME + ".net.xml")
    rou = os.path.join(HERE, NAME + ".rou.xml")
    subprocess.check_call([os.path.join(REF, "netgenerate"), "--grid", "--grid.number=10", "--grid.length=200",
                           "--default-junction-type", "traffic_light", "-o", net, "--no-warnings"],
                          stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    txt = open(net).read()
    a, b = txt.find("<!--"), txt.find("-->")
    if 0 <= a < b:
        txt = txt[:a] + txt[b + 3:]
    open(net, "w").write(txt)
    edges, succ = read_graph(net)
    rng = random.Random(SEED)
    departs = sorted(rng.uniform(0, DEPART_END) for _ in range(N_VEH))
    with open(rou, "w") as f:
        f.write('<routes>\n    <vType id="car" sigma="0.5"/>\n')
        for i, d in enumerate(departs):
            while True:
                e = rng.choice(edges)
                route = [e]
                while len(route) < ROUTE_EDGES and succ.get(route[-1]):
                    route.append(rng.choice(succ[route[-1]]))
                if len(route) == ROUTE_EDGES:
                    break
            f.write('    <vehicle id="v%d" type="car" depart="%.2f"><route edges="%s"/></vehicle>\n' % (i, d, " ".join(route)))
        f.write("</routes>\n")
    flat = os.path.join(HERE, NAME + ".flat")
    subprocess.check_call([os.path.join(REF, "sumo_flatten"), "--flat-out", flat, "-n", net, "-r", rou, "--no-step-log",
                           "--no-warnings", "--time-to-teleport", "-1", "--default.departspeed", "0", "--default.departlane", "first",
                           # every vehicle carries MSDevice_Emissions (default class HBEFA4/PC_petrol_Euro-4); bench.py switches the
                           # device off (veh_emissions = 0) unless it is run with --emissions (BASELINE.json configs[4])
                           "--device.emissions.probability", "1"],
                          stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    for p in (net, rou, flat):
        print(p, os.path.getsize(p))


if __name__ == "__main__":
    main()
