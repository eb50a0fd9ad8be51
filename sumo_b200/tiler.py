"""Replicates a flattened scenario (flatfile.load_flat) into K disjoint districts, directly in array space.

This is the large-input path of SURVEY.md section 8 row f3: a 1M-vehicle network cannot be pushed through the reference's
single-threaded SAX loader in reasonable time, so the base tile is loaded ONCE by the reference's own NLBuilder
(integration/sumo_flatten.cpp) and the engine's input is built from it by index arithmetic.  Districts do not interact,
so every vehicle-update costs exactly what it costs in the base tile; departure times are shifted per district so the
districts do not run in lock-step.
"""
import numpy as np

# arrays whose VALUES are indices into another table: name -> table
_INDEX = {
    "lane_edge": "edge", "lane_pred_logical": "lane", "lane_pred_normal": "lane", "lane_entry_link": "link",
    "lane_inc_link": "link", "inc_lane": "lane", "inc_link": "link", "lane_to_junction": "junction",
    "lane_from_junction": "junction", "link_from": "lane", "link_to": "lane", "link_via": "lane", "link_tl": "tl",
    "link_internal_before": "lane", "link_junction": "junction", "link_entry": "link", "link_exit": "link",
    "foelink": "link", "foelane": "lane", "tllink_link": "link", "route_edges": "edge", "veh_route": "route",
    "lane_numid": "lane", "veh_numid": "veh", "veh_idrank": "veh",
}
# CSR offset arrays (length n+1): name -> table the offsets point into
_CSR = {
    "lane_link_begin": "link", "lane_inc_begin": "inc", "lane_shape_begin": "shape", "edge_lane_begin": "lane",
    "link_foelink_begin": "foelink", "link_foelane_begin": "foelane", "tl_phase_begin": "phase",
    "phase_state_begin": "state", "tl_link_begin": "tllink", "route_begin": "route_edge",
}
_TABLE_SIZE_FROM = {
    "lane": "lane_length", "edge": "edge_flags", "link": "link_from", "junction": None, "tl": "tl_cur_phase",
    "route": None, "veh": "veh_type", "inc": "inc_lane", "shape": None, "foelink": "foelink", "foelane": "foelane",
    "phase": "phase_dur", "state": "phase_states", "tllink": "tllink_index", "route_edge": "route_edges",
}
_SHARED_PREFIXES = ("vt_", "opt_")
_DROP_SUFFIXES = ("_names", "_name_begin")


def _table_sizes(base):
    n = {}
    for table, src in _TABLE_SIZE_FROM.items():
        if src is not None:
            n[table] = int(base[src].size)
    n["junction"] = int(max(base["lane_to_junction"].max(initial=-1), base["lane_from_junction"].max(initial=-1),
                            base["link_junction"].max(initial=-1))) + 1
    n["route"] = int(base["route_begin"].size) - 1
    n["shape"] = int(base["shape_xy"].size) // 2
    return n


def tile(base, k, depart_shift_ms=1000, depart_shift_mod=16, overrides=None):
    """returns the arrays of a scenario made of k disjoint copies of `base`.
    overrides: {array name: new value array} applied to the shared (vt_*/opt_*) arrays, e.g. {"vt_sigma": [0.0]}"""
    n = _table_sizes(base)
    out = {}
    for name, a in base.items():
        if name.endswith(_DROP_SUFFIXES):
            continue   # ids are not replicated (diagnostics only)
        if name.startswith(_SHARED_PREFIXES):
            out[name] = a.copy()
            continue
        if name in _CSR:
            step = n[_CSR[name]]
            body = a[:-1].astype(np.int64)
            rep = (body[None, :] + (np.arange(k, dtype=np.int64) * step)[:, None]).reshape(-1)
            out[name] = np.concatenate([rep, [k * step]]).astype(a.dtype)
        elif name in _INDEX:
            step = n[_INDEX[name]]
            a64 = a.astype(np.int64)
            rep = a64[None, :] + (np.arange(k, dtype=np.int64) * step)[:, None]
            rep = np.where(a64[None, :] < 0, a64[None, :], rep)   # -1 stays "none"
            out[name] = rep.reshape(-1).astype(a.dtype)
        elif name == "veh_depart":
            shift = (np.arange(k, dtype=np.int64) % depart_shift_mod) * depart_shift_ms
            out[name] = (a[None, :] + shift[:, None]).reshape(-1)
        elif name == "link_response":
            out[name] = np.tile(a, k)
        else:
            out[name] = np.tile(a, k)
    if overrides:
        for name, v in overrides.items():
            out[name] = np.asarray(v, dtype=out[name].dtype).reshape(out[name].shape)
    return out
