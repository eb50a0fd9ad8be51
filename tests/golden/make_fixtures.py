#!/usr/bin/env python3
"""Generates the committed scenario fixtures under tests/golden/ (run in the build container, needs oracle/_ref).

  netgenerate (the UNMODIFIED reference binary, oracle/_ref/netgenerate)  ->  <name>.net.xml
  random trips (python random, seeded; shortest hop paths on the connection graph)  ->  <name>.rou.xml

The GPU box has no /root/reference, so the fixtures are committed; oracle/_ref/sumo_flatten (which travels with the
snapshot) turns them into .flat + reference traces at test time.

usage: python tests/golden/make_fixtures.py [name ...]
"""
import os
import random
import subprocess
import sys
import xml.etree.ElementTree as ET
from collections import deque

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "..", "..", "oracle", "_ref")


def read_graph(net_path):
    """normal edges and the edge->edge connection graph of a .net.xml"""
    edges = []
    succ = {}
    for _, el in ET.iterparse(net_path):
        if el.tag == "edge":
            if el.get("function") is None:
                edges.append(el.get("id"))
            el.clear()
        elif el.tag == "connection":
            f, t = el.get("from"), el.get("to")
            if not f.startswith(":") and not t.startswith(":") and el.get("dir") != "t":
                succ.setdefault(f, [])
                if t not in succ[f]:
                    succ[f].append(t)
            el.clear()
    return edges, succ


def shortest(succ, a, b):
    prev = {a: None}
    q = deque([a])
    while q:
        x = q.popleft()
        if x == b:
            break
        for y in succ.get(x, []):
            if y not in prev:
                prev[y] = x
                q.append(y)
    if b not in prev:
        return None
    path = []
    x = b
    while x is not None:
        path.append(x)
        x = prev[x]
    return path[::-1]


def make_routes(net_path, out_path, n_veh, period_end, seed, vtype_attrs, min_edges=3):
    edges, succ = read_graph(net_path)
    rng = random.Random(seed)
    departs = sorted(rng.uniform(0, period_end) for _ in range(n_veh))
    with open(out_path, "w") as f:
        f.write("<routes>\n")
        f.write('    <vType id="car" %s/>\n' % vtype_attrs)
        i = 0
        for d in departs:
            while True:
                a, b = rng.choice(edges), rng.choice(edges)
                if a == b:
                    continue
                p = shortest(succ, a, b)
                if p is not None and len(p) >= min_edges:
                    break
            f.write('    <vehicle id="v%d" type="car" depart="%.2f"><route edges="%s"/></vehicle>\n' % (i, d, " ".join(p)))
            i += 1
        f.write("</routes>\n")


SCENARIOS = {
    # config 1 of BASELINE.json: 10x10 grid, 400 m, priority junctions, ~2000 vehicles / 3600 s, Krauss sigma=0
    "c1_grid10": dict(netgen=["--grid", "--grid.number=10", "--grid.length=400"], n=2000, end=3600, seed=42,
                      vtype='sigma="0"'),
    # the same net under heavy load (queues, junction blocking, impatience) for a shorter horizon
    "grid5_dense": dict(netgen=["--grid", "--grid.number=5", "--grid.length=150"], n=1500, end=600, seed=7,
                        vtype='sigma="0"'),
    # static traffic lights (config 2 in miniature)
    "grid4_tls": dict(netgen=["--grid", "--grid.number=4", "--grid.length=200", "--default-junction-type", "traffic_light"],
                      n=800, end=900, seed=11, vtype='sigma="0"'),
    # other network shapes / junction types of netgenerate
    "rand_net": dict(netgen=["--rand", "--rand.iterations=40", "--seed", "5"], n=500, end=400, seed=3, vtype='sigma="0"', min_edges=2),
    "grid4_stop": dict(netgen=["--grid", "--grid.number=4", "--grid.length=150", "--default-junction-type", "priority_stop"],
                       n=500, end=400, seed=3, vtype='sigma="0"', min_edges=2),
    "grid3_tls_short": dict(netgen=["--grid", "--grid.number=3", "--grid.length=60", "--default-junction-type", "traffic_light"],
                            n=500, end=400, seed=3, vtype='sigma="0"', min_edges=2),
    # a tiny case that finishes in a blink (smoke)
    "grid3_small": dict(netgen=["--grid", "--grid.number=3", "--grid.length=100"], n=60, end=120, seed=3, vtype='sigma="0"'),
}


def main():
    names = sys.argv[1:] or list(SCENARIOS)
    for name in names:
        sc = SCENARIOS[name]
        net = os.path.join(HERE, name + ".net.xml")
        rou = os.path.join(HERE, name + ".rou.xml")
        subprocess.check_call([os.path.join(REF, "netgenerate")] + sc["netgen"] + ["-o", net, "--no-warnings"],
                              stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        # drop the generated-on timestamp comment so the fixture is reproducible
        txt = open(net).read()
        start = txt.find("<!--")
        stop = txt.find("-->", start)
        if start >= 0 and stop > start:
            txt = txt[:start] + txt[stop + 3:]
        open(net, "w").write(txt)
        make_routes(net, rou, sc["n"], sc["end"], sc["seed"], sc["vtype"], sc.get("min_edges", 3))
        print(name, os.path.getsize(net), os.path.getsize(rou))


if __name__ == "__main__":
    main()
