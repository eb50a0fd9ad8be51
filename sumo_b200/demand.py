"""Bulk demand ingestion: the vehicles and routes of a .rou.xml as the veh_* / route_* arrays of
include/sumob200_flatfile.h, without the reference's per-vehicle host objects (MSRouteHandler::closeVehicle,
src/microsim/MSRouteHandler.cpp:605 -> MSVehicleControl::buildVehicle, src/microsim/MSVehicleControl.cpp:109: one MSVehicle, one
std::map node and one MSRoute per vehicle -- what makes a 10 M vehicle scenario a multi-GB host heap before the first step).

Scope: <vehicle id type depart [departPos departSpeed departLane arrivalPos]> with an embedded <route edges="..."/> or a route=
reference, <route id edges>, vTypes by id (their parameters come from the network / type arrays already flattened).  The order
rules are the reference's: vehicles are inserted in the order of MSVehicleContainer (depart time rounded up to the step grid,
then load order); numerical ids count loaded vehicles; veh_idrank is the std::string order of the ids (MSVehicle::isLeader's tie
breaker, FCD output order).  Validated against the arrays the reference's own loader produces (tests/test_demand.py).

gridgen.make_demand generates synthetic demand straight into the same arrays."""
import re

import numpy as np

from .flatfile import names as flat_names

_PROC_POS = {"random": 3, "free": 4, "base": 5, "last": 6, "random_free": 7}
_PROC_SPEED = {"random": 3, "max": 4, "desired": 5, "speedLimit": 6, "last": 7, "avg": 8}
_PROC_LANE = {"random": 2, "free": 3, "allowed": 4, "best": 5, "first": 7}


def _attrs(text):
    return dict(re.findall(r'([\w.:]+)="([^"]*)"', text))


def demand_from_routes_xml(net, rou_path, delta_t_ms=1000):
    """net: arrays of the flattened network incl. edge names and vt_names.  Returns the dict of route_* / veh_* arrays."""
    edge_index = {s: i for i, s in enumerate(flat_names(net, "edge"))}
    type_index = {s: i for i, s in enumerate(flat_names(net, "vt"))}
    text = open(rou_path).read()
    named_routes = {}
    for m in re.finditer(r"<route\s+([^>]*?)/?>", text):
        a = _attrs(m.group(1))
        if "id" in a:
            named_routes[a["id"]] = a["edges"].split()
    vehicles = []
    for m in re.finditer(r"<vehicle\s+([^>]*?)(/>|>(.*?)</vehicle>)", text, re.S):
        a = _attrs(m.group(1))
        body = m.group(3) or ""
        if "<stop" in body:
            raise ValueError("vehicle '%s': <stop>s are not supported" % a["id"])
        if "route" in a:
            edges = named_routes[a["route"]]
        else:
            edges = _attrs(re.search(r"<route\s+([^>]*?)/?>", body).group(1))["edges"].split()
        vehicles.append((a, edges))
    n = len(vehicles)
    depart = np.array([int(round(float(a["depart"]) * 1000.)) for a, _ in vehicles], np.int64)
    bucket = depart + (-depart) % delta_t_ms
    order = np.lexsort((np.arange(n), bucket))          # MSVehicleContainer: depart bucket, then load order (= numerical id)
    route_begin = np.zeros(n + 1, np.int32)
    route_edges = []
    out = {k: np.zeros(n, np.int32) for k in ("veh_type", "veh_route", "veh_numid", "veh_departpos_proc", "veh_departspeed_proc",
                                              "veh_departlane_proc", "veh_departlane", "veh_idrank")}
    out.update({k: np.zeros(n, np.float64) for k in ("veh_departpos", "veh_arrivalpos", "veh_speedfactor")})
    out["veh_departspeed"] = -np.ones(n)
    out["veh_depart"] = depart[order]
    out["veh_emissions"] = np.zeros(n, np.uint8)
    ids = []
    for slot, src in enumerate(order):
        a, edges = vehicles[src]
        ids.append(a["id"])
        e = [edge_index[x] for x in edges]
        route_edges += e
        route_begin[slot + 1] = len(route_edges)
        out["veh_route"][slot] = slot
        out["veh_numid"][slot] = src
        t = type_index[a.get("type", "DEFAULT_VEHTYPE")]
        out["veh_type"][slot] = t
        out["veh_speedfactor"][slot] = 1.0      # speedFactor deviation 0 (--default.speeddev 0); sampled values come from the caller
        last_lane = net["edge_lane_begin"][e[-1]]
        out["veh_arrivalpos"][slot] = float(a["arrivalPos"]) if "arrivalPos" in a else net["lane_length"][last_lane]
        if "departPos" in a:
            v = a["departPos"]
            out["veh_departpos_proc"][slot] = _PROC_POS.get(v, 1)
            out["veh_departpos"][slot] = 0. if v in _PROC_POS else float(v)
        if "departSpeed" in a:
            v = a["departSpeed"]
            out["veh_departspeed_proc"][slot] = _PROC_SPEED.get(v, 1)
            out["veh_departspeed"][slot] = -1. if v in _PROC_SPEED else float(v)
        if "departLane" in a:
            v = a["departLane"]
            out["veh_departlane_proc"][slot] = _PROC_LANE.get(v, 1)
            out["veh_departlane"][slot] = 0 if v in _PROC_LANE else int(v)
    rank = np.zeros(n, np.int32)
    rank[np.argsort(np.array(ids, dtype=object), kind="stable")] = np.arange(n, dtype=np.int32)
    out["veh_idrank"] = rank
    out["route_begin"] = route_begin
    out["route_edges"] = np.array(route_edges, np.int32)
    begin = np.zeros(n + 1, np.int32)
    raw = bytearray()
    for i, s in enumerate(ids):
        raw += s.encode()
        begin[i + 1] = len(raw)
    out["veh_name_begin"], out["veh_names"] = begin, np.frombuffer(bytes(raw), np.uint8).copy()
    return out
