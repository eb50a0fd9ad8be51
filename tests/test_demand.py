"""Bulk demand ingestion (sumo_b200/demand.py): the veh_* / route_* arrays built straight from a .rou.xml must be
the arrays the reference's own loader (MSRouteHandler -> MSVehicleControl::buildVehicle, flattened by oracle/_ref/sumo_flatten)
produces, and the step logic run on them must reproduce the reference's trajectories."""
import subprocess

import numpy as np
import pytest

from conftest import GOLDEN, reference_case
from sumo_b200.demand import demand_from_routes_xml
from sumo_b200.flatfile import load_flat, load_trace, save_flat
from test_parity_host import assert_traces_identical
import os


@pytest.mark.parametrize("case", ["c1_grid10", "grid4_tls", "grid5_dense"])
def test_demand_arrays_equal_the_loaders(case, host_emul, tmp_path):
    flat, trace = reference_case(case)
    ref = load_flat(flat)
    net = {k: v for k, v in ref.items() if not (k.startswith("veh_") or k.startswith("route_"))}
    d = demand_from_routes_xml(net, os.path.join(GOLDEN, case + ".rou.xml"))
    # routes: the loader shares identical routes between vehicles only by id; compare per vehicle
    for k in ("veh_depart", "veh_type", "veh_numid", "veh_idrank", "veh_arrivalpos", "veh_speedfactor", "veh_departpos_proc", "veh_departspeed_proc",
              "veh_departlane_proc"):
        assert np.array_equal(d[k], ref[k]), k
    for v in range(len(d["veh_type"])):
        a = d["route_edges"][d["route_begin"][d["veh_route"][v]]:d["route_begin"][d["veh_route"][v] + 1]]
        b = ref["route_edges"][ref["route_begin"][ref["veh_route"][v]]:ref["route_begin"][ref["veh_route"][v] + 1]]
        assert np.array_equal(a, b)
    net.update(d)
    p = str(tmp_path / "x.flat")
    save_flat(p, net)
    out = str(tmp_path / "x.trace")
    reft = load_trace(trace)[:1200]
    subprocess.run([host_emul, p, "--trace-out", out, "--steps", str(len(reft))], check=True, timeout=600, stderr=subprocess.DEVNULL)
    assert assert_traces_identical(reft, load_trace(out)) > 1000
