"""sumo_b200/gridgen.py: the directly generated N x N grid against what the reference's own netgenerate +
loader produce for the same N (oracle/_ref/netgenerate, oracle/_ref/sumo_flatten), array by array.  Identical: every lane / link
attribute, every cross reference (successor links, via lanes, entry / exit links, incoming lanes, foe links), the right-of-way
response matrices, the signal programs, the lane shapes, the sets of foe lanes with their conflict flags.  Equal to 1e-9: the
three quantities the loader derives from ABSOLUTE coordinates (lengths behind crossing / conflict sizes, link radius, lane
geometry factor: they differ in the last bits between two junctions of the same network).  Order inside a link's foe-lane list:
netbuild's own order varies between equal junctions of one network (71 of 1984 links on 8 x 8), compared as sets.

Sizes up to 10 x 10: netbuild breaks right-of-way ties between equal-priority streams by the STRING order of the edge ids
(src/netbuild/NBRequest.cpp), so in an 11 x 11 grid the junctions next to row "10" ("B10B9" < "B8B9") get a mirrored response
matrix for two of their links.  The generator keeps, for every N, the matrix netgenerate gives a junction whose ids sort in numeric
order -- a valid right of way, and the one every junction of a <= 10 x 10 grid has."""
import os
import subprocess

import numpy as np
import pytest

from conftest import CACHE, GOLDEN, REF, ensure_reference
from sumo_b200 import gridgen
from sumo_b200.flatfile import load_flat, names


def reference_grid(n, length, jtype):
    ensure_reference()
    os.makedirs(CACHE, exist_ok=True)
    base = os.path.join(CACHE, "refgrid_%d_%d_%s" % (n, length, jtype))
    if not os.path.exists(base + ".flat"):
        subprocess.check_call([os.path.join(REF, "netgenerate"), "--grid", "--grid.number=%d" % n, "--grid.length=%d" % length,
                               "--default-junction-type", jtype, "-o", base + ".net.xml", "--no-warnings"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        with open(base + ".rou.xml", "w") as f:
            f.write('<routes><vType id="car" sigma="0.5"/><vehicle id="v0" type="car" depart="0"><route edges="A0A1 A1A2"/></vehicle></routes>\n')
        subprocess.check_call([os.path.join(REF, "sumo_flatten"), "--flat-out", base + ".flat", "-n", base + ".net.xml", "-r", base + ".rou.xml",
                               "--no-step-log", "--no-warnings", "--default.departspeed", "0", "--default.departlane", "first"],
                              stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return load_flat(base + ".flat")


def rows(begin, payload, idx, scale=1):
    return [payload[begin[i] * scale:begin[i + 1] * scale] for i in idx]


@pytest.mark.parametrize("template,n,length,jtype", [("grid5_tls800", 8, 800, "traffic_light"), ("grid5_tls800", 10, 800, "traffic_light"),
                                                    ("grid5_prio100", 9, 100, "priority")])
def test_generated_grid_equals_netgenerate(template, n, length, jtype):
    t = load_flat(os.path.join(GOLDEN, template + ".flat"))
    r = reference_grid(n, length, jtype)
    g = gridgen.make_grid(t, n)
    assert float(g["grid_spacing"][0]) == float(length)
    gl, rl = names(g, "lane"), names(r, "lane")
    assert sorted(gl) == sorted(rl)
    ridx = {s: i for i, s in enumerate(rl)}
    perm = np.array([ridx[s] for s in gl])          # generated lane l is reference lane perm[l]
    inv = np.zeros(len(perm), np.int64)
    inv[perm] = np.arange(len(perm))

    def lane_ref(a):
        return np.where(a >= 0, inv[np.where(a >= 0, a, 0)], -1)
    for k in ("lane_length", "lane_speed", "lane_flags", "lane_width"):
        assert np.array_equal(g[k], r[k][perm]), k
    assert np.abs(g["lane_geomfactor"] - r["lane_geomfactor"][perm]).max() < 1e-9
    for k in ("lane_pred_logical", "lane_pred_normal"):
        assert np.array_equal(g[k], lane_ref(r[k][perm])), k
    lperm = np.concatenate([np.arange(r["lane_link_begin"][p], r["lane_link_begin"][p + 1]) for p in perm])
    assert np.array_equal(np.diff(g["lane_link_begin"]), np.diff(r["lane_link_begin"])[perm])
    linv = np.zeros(len(lperm), np.int64)
    linv[lperm] = np.arange(len(lperm))

    def link_ref(a):
        return np.where(a >= 0, linv[np.where(a >= 0, a, 0)], -1)
    for k in ("link_state", "link_dir", "link_length", "link_foevis", "link_flags", "link_index", "link_tlindex"):
        assert np.array_equal(g[k], r[k][lperm]), k
    finite = g["link_radius"] < 1e300
    assert np.array_equal(finite, r["link_radius"][lperm] < 1e300)
    assert np.abs(g["link_radius"][finite] - r["link_radius"][lperm][finite]).max() < 1e-9
    for k in ("link_from", "link_to", "link_via", "link_internal_before"):
        assert np.array_equal(g[k], lane_ref(r[k][lperm])), k
    for k in ("link_entry", "link_exit"):
        assert np.array_equal(g[k], link_ref(r[k][lperm])), k
    assert np.array_equal(g["link_response"].reshape(-1, 4), r["link_response"].reshape(-1, 4)[lperm])
    assert np.array_equal(g["lane_entry_link"], link_ref(r["lane_entry_link"][perm]))
    nk = len(lperm)
    for a, b in zip(rows(g["link_foelink_begin"], g["foelink"], range(nk)), rows(r["link_foelink_begin"], link_ref(r["foelink"]), lperm)):
        assert np.array_equal(a, b)
    # foe lanes: same set per link; conflict records matched by foe lane
    rfl = lane_ref(r["foelane"])
    n_conf = n_touch = 0
    for k in range(nk):
        ga, gb = g["link_foelane_begin"][k], g["link_foelane_begin"][k + 1]
        ra, rb = r["link_foelane_begin"][lperm[k]], r["link_foelane_begin"][lperm[k] + 1]
        assert gb - ga == rb - ra
        go = np.argsort(g["foelane"][ga:gb], kind="stable")
        ro = np.argsort(rfl[ra:rb], kind="stable")
        assert np.array_equal(g["foelane"][ga:gb][go], rfl[ra:rb][ro])
        same = g["conf_flag"][ga:gb][go] == r["conf_flag"][ra:rb][ro]
        n_conf += gb - ga
        # the loader's polyline intersection test (MSLink::setRequestInformation, MSLink.cpp:259-420) is not translation invariant
        # for internal lanes that only touch: the same pair of lanes gets CONFLICT_NO_INTERSECTION at one junction and a crossing
        # with a length of a few centimetres at another junction of the same network.  Counted, must stay a tiny share.
        n_touch += int((~same).sum())
        assert all(2 in (a, b) for a, b in zip(g["conf_flag"][ga:gb][go][~same], r["conf_flag"][ra:rb][ro][~same]))
        for c in ("conf_lbc", "conf_size"):
            assert np.abs(g[c][ga:gb][go][same] - r[c][ra:rb][ro][same]).max(initial=0) < 1e-9
    assert n_touch <= 0.01 * n_conf, (n_touch, n_conf)
    nl = len(perm)
    for a, b in zip(rows(g["lane_inc_begin"], g["inc_lane"], range(nl)), rows(r["lane_inc_begin"], lane_ref(r["inc_lane"]), perm)):
        assert np.array_equal(a, b)
    for a, b in zip(rows(g["lane_inc_begin"], g["inc_length"], range(nl)), rows(r["lane_inc_begin"], r["inc_length"], perm)):
        assert np.array_equal(a, b)
    for a, b in zip(rows(g["lane_shape_begin"], g["shape_xy"], range(nl), 2), rows(r["lane_shape_begin"], r["shape_xy"], perm, 2)):
        assert np.array_equal(a, b)
    gt, rt = names(g, "tl"), names(r, "tl")
    assert sorted(gt) == sorted(rt)
    for i, s in enumerate(gt):
        j = rt.index(s)
        assert np.array_equal(g["phase_dur"][g["tl_phase_begin"][i]:g["tl_phase_begin"][i + 1]], r["phase_dur"][r["tl_phase_begin"][j]:r["tl_phase_begin"][j + 1]])
        assert g["tl_next_switch"][i] == r["tl_next_switch"][j] and g["tl_cur_phase"][i] == r["tl_cur_phase"][j]
        gs = g["phase_states"][g["phase_state_begin"][g["tl_phase_begin"][i]]:g["phase_state_begin"][g["tl_phase_begin"][i + 1]]]
        rs = r["phase_states"][r["phase_state_begin"][r["tl_phase_begin"][j]]:r["phase_state_begin"][r["tl_phase_begin"][j + 1]]]
        assert np.array_equal(gs, rs)
        assert np.array_equal(g["tllink_link"][g["tl_link_begin"][i]:g["tl_link_begin"][i + 1]], link_ref(r["tllink_link"][r["tl_link_begin"][j]:r["tl_link_begin"][j + 1]]))


def test_tls_offsets_follow_the_loader(tmp_path):
    """set_tls_offsets against the reference's loader reading <tlLogic offset=...> (NLJunctionControlBuilder.cpp:238-252)"""
    import re
    ensure_reference()
    reference_grid(8, 800, "traffic_light")
    net = open(os.path.join(CACHE, "refgrid_8_800_traffic_light.net.xml")).read()
    offsets = {}

    def repl(m):
        off = (len(offsets) * 37) % 90 - 20     # positive and negative offsets
        offsets[m.group(1)] = off
        return '<tlLogic id="%s" type="static" programID="0" offset="%d">' % (m.group(1), off)
    net = re.sub(r'<tlLogic id="([^"]+)" type="static" programID="0" offset="0">', repl, net)
    assert len(offsets) == 64
    p = str(tmp_path / "off.net.xml")
    open(p, "w").write(net)
    subprocess.check_call([os.path.join(REF, "sumo_flatten"), "--flat-out", str(tmp_path / "off.flat"), "-n", p, "-r", os.path.join(CACHE, "refgrid_8_800_traffic_light.rou.xml"),
                           "--no-step-log", "--no-warnings"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    r = load_flat(str(tmp_path / "off.flat"))
    g = gridgen.make_grid(load_flat(os.path.join(GOLDEN, "grid5_tls800.flat")), 8)
    gt, rt = names(g, "tl"), names(r, "tl")
    gridgen.set_tls_offsets(g, [offsets[s] * 1000 for s in gt])
    for i, s in enumerate(gt):
        j = rt.index(s)
        assert (g["tl_cur_phase"][i], g["tl_next_switch"][i]) == (r["tl_cur_phase"][j], r["tl_next_switch"][j]), s


def test_demand_is_drivable(host_emul, tmp_path):
    """a generated grid with generated demand runs in the step logic without a single error flag, every vehicle departs"""
    from sumo_b200.flatfile import save_flat
    g = gridgen.make_grid(load_flat(os.path.join(GOLDEN, "grid5_tls800.flat")), 7)
    gridgen.set_tls_offsets(g, (np.arange(49) * 13 % 90) * 1000)
    gridgen.make_demand(g, 2500, 10, 120, seed=3, sigma=0.5, speeddev=0.1)
    p = str(tmp_path / "g7.flat")
    save_flat(p, g)
    out = subprocess.run([host_emul, p, "--steps", "300"], check=True, capture_output=True, text=True, timeout=600).stderr
    assert "inserted=2500" in out and "err=0x0" in out
    # ranks of "v<i>" ids in std::string order
    ids = ["v%d" % i for i in range(1234)]
    want = np.argsort(np.argsort(np.array(ids, dtype=object), kind="stable"))
    assert np.array_equal(gridgen._decimal_string_rank(1234), want)
