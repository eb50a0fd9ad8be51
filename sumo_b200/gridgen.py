"""Direct generator of flattened N x N grid networks: what `netgenerate --grid --grid.number=N`
(src/netgen/NGNet.cpp:93 createChequerBoard) + netbuild (junction shapes, internal lanes, right-of-way matrices, traffic-light
programs: src/netbuild/NBNode.cpp, NBRequest.cpp, NBOwnTLDef.cpp) + the reference's loader (NLBuilder, MSLink::setRequestInformation)
produce, written straight into the arrays of include/sumob200_flatfile.h -- without a .net.xml, which for the 1000 x 1000 grid of
BASELINE.json configs[3]/[4] would be tens of GB pushed through a single-threaded SAX loader.

How: a chequer-board grid is translation invariant.  Every junction looks like one of 5 x 5 kinds, given by whether its column
(and row) index is 0, 1, interior, N-2 or N-1 -- its own arms and the arms of its direct neighbours decide every number
netbuild computes for it.  So ONE small network that netgenerate itself built (a 5 x 5 grid of the same spacing and options, the
"template", flattened by the reference's own loader) contains every kind once, and the N x N network is that template's
junctions, lanes, links, conflict tables and signal programs instantiated per target junction with all cross references
re-based (`_Stencil`).  Lane lengths, speeds, shapes, conflict geometry, response matrices are therefore the reference's own
numbers, not a re-derivation.  Validated against netgenerate + the reference loader on 8 x 8 / 10 x 10 grids, array by array
(tests/test_gridgen.py).

Also here: demand for such a grid (random-walk routes without U-turns, departures spread over an interval), generated as arrays
(no per-vehicle host objects), and per-junction traffic-light offsets (NLJunctionControlBuilder.cpp:238-252).
"""
import numpy as np

from .flatfile import names as flat_names

LF_INTERNAL = 1


def _pack_names(strings):
    begin = np.zeros(len(strings) + 1, np.int32)
    raw = bytearray()
    for i, s in enumerate(strings):
        raw += s.encode()
        begin[i + 1] = len(raw)
    return begin, np.frombuffer(bytes(raw), np.uint8).copy()


def alphabetical_code(i, imax):
    """NGNet::alphabeticalCode (src/netgen/NGNet.cpp:80-90)"""
    xn = 1
    while 26 ** xn < imax:
        xn += 1
    out = ""
    for _ in range(xn):
        out = chr(ord("A") + i % 26) + out
        i //= 26
    return out


def _csr_gather(begin, idx):
    """for rows idx of a CSR structure: (new begin array, flat positions into the payload)"""
    cnt = begin[idx + 1] - begin[idx]
    nb = np.zeros(len(idx) + 1, np.int64)
    np.cumsum(cnt, out=nb[1:])
    pos = np.repeat(begin[idx] - nb[:-1], cnt) + np.arange(nb[-1])
    return nb, pos, cnt


class _Stencil:
    """the template network, indexed for instantiation"""

    def __init__(self, t):
        self.t = t
        jn = flat_names(t, "junction")
        self.T = int(round(len(jn) ** 0.5))
        T = self.T
        if T * T != len(jn) or T < 5:
            raise ValueError("the template must be a square netgenerate grid of at least 5 x 5 junctions")
        code = {alphabetical_code(i, T): i for i in range(T)}
        self.jxy = np.zeros((len(jn), 2), np.int64)
        for j, name in enumerate(jn):
            k = 0
            while not name[k].isdigit():
                k += 1
            self.jxy[j] = (code[name[:k]], int(name[k:]))
        self.jat = -np.ones((T, T), np.int64)
        self.jat[self.jxy[:, 0], self.jxy[:, 1]] = np.arange(len(jn))
        nL = len(t["lane_length"])
        if len(t["edge_flags"]) != nL:
            raise ValueError("the template must have one lane per edge")
        self.internal = (t["lane_flags"] & LF_INTERNAL) != 0
        toj = t["lane_to_junction"].astype(np.int64)
        # per template junction: its internal lanes, then its incoming normal lanes, in template order = the junction's lane block
        order = np.lexsort((np.arange(nL), ~self.internal, toj))
        self.block_lanes = order
        self.block_begin = np.searchsorted(toj[order], np.arange(len(jn) + 1))
        self.local = np.zeros(nL, np.int64)        # position of a template lane inside its junction's block
        self.local[order] = np.arange(nL) - self.block_begin[toj[order]]
        self.lane_owner = toj
        # traffic lights: one program per junction, found through the links it controls
        self.tl_of_junction = -np.ones(len(jn), np.int64)
        k = np.nonzero(t["link_tl"] >= 0)[0]
        self.tl_of_junction[t["link_junction"][k]] = t["link_tl"][k]


def _kind(i, n, T):
    """template row/column index that has the same surroundings as index i of n"""
    if i == 0:
        return 0
    if i == 1:
        return 1
    if i == n - 1:
        return T - 1
    if i == n - 2:
        return T - 2
    return 2


def make_grid(template, n, spacing=None, with_names=True):
    """the arrays of an n x n grid with the template's spacing, junction type, lane count and speed.  Returns a dict of arrays
    like flatfile.load_flat (network part + options; add demand with make_demand)."""
    st = _Stencil(template)
    t = template
    T = st.T
    if n < 5:
        raise ValueError("n must be >= 5")
    kinds = np.array([_kind(i, n, T) for i in range(n)], np.int64)
    # target junction index = ix * n + iy (netgenerate creates the nodes column by column, NGNet.cpp:95-99)
    ix, iy = np.divmod(np.arange(n * n, dtype=np.int64), n)
    tj = st.jat[kinds[ix], kinds[iy]]                       # template junction instantiated at each target junction
    nb = st.block_begin[tj + 1] - st.block_begin[tj]        # lanes per target junction block
    lane_base = np.zeros(n * n + 1, np.int64)
    np.cumsum(nb, out=lane_base[1:])
    nL = int(lane_base[-1])
    lane_j = np.repeat(np.arange(n * n, dtype=np.int64), nb)            # target junction owning each target lane
    lane_t = st.block_lanes[np.repeat(st.block_begin[tj] - lane_base[:-1], nb) + np.arange(nL)]   # template lane of each target lane

    if spacing is None:
        # distance between neighbouring junction centres, from the template's straight boundary lane A0 -> A1 ... robustly: shapes
        a, b = st.jat[1, 1], st.jat[2, 1]
        spacing = float(_junction_pos(t, st, b)[0] - _junction_pos(t, st, a)[0])

    def map_junction(anchor_j, tmpl_j):
        """target junction for template junction tmpl_j seen from an entity instantiated at target junction anchor_j"""
        d = st.jxy[tmpl_j] - st.jxy[tj[anchor_j]]
        return (ix[anchor_j] + d[:, 0]) * n + (iy[anchor_j] + d[:, 1])

    def map_lane(anchor_j, tmpl_lane):
        """target lane for template lane ids referenced by entities instantiated at target junctions anchor_j (-1 stays -1)"""
        tmpl_lane = np.asarray(tmpl_lane, np.int64)
        ok = tmpl_lane >= 0
        x = np.where(ok, tmpl_lane, 0)
        out = lane_base[map_junction(anchor_j, st.lane_owner[x])] + st.local[x]
        return np.where(ok, out, -1)

    g = {}
    # ---- options and everything that is not network: copied
    for k, v in t.items():
        if k.startswith("opt_") or k.startswith("vt_"):
            g[k] = v.copy()
    # ---- lanes: plain attributes
    for k in ("lane_length", "lane_speed", "lane_width", "lane_flags", "lane_geomfactor"):
        g[k] = t[k][lane_t].copy()
    g["lane_numid"] = np.arange(nL, dtype=np.int32)
    g["lane_edge"] = np.arange(nL, dtype=np.int32)
    g["edge_flags"] = t["edge_flags"][t["lane_edge"][lane_t]].copy()
    g["edge_lane_begin"] = np.arange(nL + 1, dtype=np.int32)
    for k in ("lane_to_junction", "lane_from_junction"):
        g[k] = map_junction(lane_j, t[k][lane_t].astype(np.int64)).astype(np.int32)
    for k in ("lane_pred_logical", "lane_pred_normal"):
        g[k] = map_lane(lane_j, t[k][lane_t]).astype(np.int32)
    # ---- links (grouped by from-lane)
    lb, lpos, lcnt = _csr_gather(t["lane_link_begin"].astype(np.int64), lane_t)
    nK = int(lb[-1])
    g["lane_link_begin"] = lb.astype(np.int32)
    link_t = lpos                                             # template link of each target link
    link_lane = np.repeat(np.arange(nL, dtype=np.int64), lcnt)  # target from-lane of each target link
    link_j = lane_j[link_lane]

    def map_link(anchor_j, tmpl_link):
        tmpl_link = np.asarray(tmpl_link, np.int64)
        ok = tmpl_link >= 0
        x = np.where(ok, tmpl_link, 0)
        fl = t["link_from"][x].astype(np.int64)
        out = lb[map_lane(anchor_j, fl)] + (x - t["lane_link_begin"][fl])
        return np.where(ok, out, -1)

    for k in ("link_state", "link_dir", "link_length", "link_foevis", "link_radius", "link_flags", "link_index", "link_tlindex"):
        g[k] = t[k][link_t].copy()
    for k in ("link_from", "link_to", "link_via", "link_internal_before"):
        g[k] = map_lane(link_j, t[k][link_t]).astype(np.int32)
    for k in ("link_entry", "link_exit"):
        g[k] = map_link(link_j, t[k][link_t]).astype(np.int32)
    g["link_junction"] = map_junction(link_j, t["link_junction"][link_t].astype(np.int64)).astype(np.int32)
    g["link_response"] = t["link_response"].reshape(-1, 4)[link_t].reshape(-1).copy()
    for k in ("lane_entry_link", "lane_inc_link"):
        g[k] = map_link(lane_j, t[k][lane_t]).astype(np.int32)
    # incoming lanes of every lane
    ib, ipos, icnt = _csr_gather(t["lane_inc_begin"].astype(np.int64), lane_t)
    g["lane_inc_begin"] = ib.astype(np.int32)
    inc_j = np.repeat(lane_j, icnt)
    g["inc_lane"] = map_lane(inc_j, t["inc_lane"][ipos]).astype(np.int32)
    g["inc_link"] = map_link(inc_j, t["inc_link"][ipos]).astype(np.int32)
    g["inc_length"] = t["inc_length"][ipos].copy()
    # foe links / foe lanes with conflict geometry
    fb, fpos, fcnt = _csr_gather(t["link_foelink_begin"].astype(np.int64), link_t)
    g["link_foelink_begin"] = fb.astype(np.int32)
    g["foelink"] = map_link(np.repeat(link_j, fcnt), t["foelink"][fpos]).astype(np.int32)
    cb, cpos, ccnt = _csr_gather(t["link_foelane_begin"].astype(np.int64), link_t)
    g["link_foelane_begin"] = cb.astype(np.int32)
    g["foelane"] = map_lane(np.repeat(link_j, ccnt), t["foelane"][cpos]).astype(np.int32)
    for k in ("conf_lbc", "conf_size", "conf_flag", "conf_foeidx"):
        g[k] = t[k][cpos].copy()
    # lane shapes: translated with the junction
    sb, spos, scnt = _csr_gather(t["lane_shape_begin"].astype(np.int64), lane_t)
    g["lane_shape_begin"] = sb.astype(np.int32)
    xy = t["shape_xy"].reshape(-1, 2)[spos].copy()
    sj = np.repeat(lane_j, scnt)
    shift = np.stack([(ix[sj] - st.jxy[tj[sj], 0]) * spacing, (iy[sj] - st.jxy[tj[sj], 1]) * spacing], axis=1)
    g["shape_xy"] = np.round(xy + shift, 2).reshape(-1)   # net.xml coordinates have 2 decimals: the nearest double, like the parser
    # ---- traffic lights: the program of the template junction, one per target junction that has one
    ttl = st.tl_of_junction[tj]
    has_tl = ttl >= 0
    tl_index = np.where(has_tl, np.cumsum(has_tl) - 1, -1)
    tl_j = np.nonzero(has_tl)[0]
    ttl = ttl[tl_j]
    g["link_tl"] = np.where(t["link_tl"][link_t] >= 0, tl_index[link_j], -1).astype(np.int32)
    pb, ppos, pcnt = _csr_gather(t["tl_phase_begin"].astype(np.int64), ttl)
    g["tl_phase_begin"] = pb.astype(np.int32)
    g["phase_dur"] = t["phase_dur"][ppos].copy()
    qb, qpos, _ = _csr_gather(t["phase_state_begin"].astype(np.int64), ppos)
    g["phase_state_begin"] = qb.astype(np.int32)
    g["phase_states"] = t["phase_states"][qpos].copy()
    kb, kpos, kcnt = _csr_gather(t["tl_link_begin"].astype(np.int64), ttl)
    g["tl_link_begin"] = kb.astype(np.int32)
    g["tllink_index"] = t["tllink_index"][kpos].copy()
    g["tllink_link"] = map_link(np.repeat(tl_j, kcnt), t["tllink_link"][kpos]).astype(np.int32)
    g["tl_static"] = t["tl_static"][ttl].copy() if "tl_static" in t else np.ones(len(ttl), np.int32)
    g["tl_cur_phase"] = t["tl_cur_phase"][ttl].copy()
    g["tl_next_switch"] = t["tl_next_switch"][ttl].copy()
    # ---- names (ids as netgenerate would give them: column letters + row number); skipped for very large grids, where the id
    # order of the lanes (MSEdgeControl's integration order, only observable through sb2_link.h veh_executes_before) falls back
    # to the index order
    if with_names:
        jname = [alphabetical_code(int(a), n) + str(int(b)) for a, b in zip(ix, iy)]
        tl_names = [jname[j] for j in tl_j]
        tln = flat_names(t, "lane")
        tjn = flat_names(t, "junction")
        suffix = [tln[x][1 + len(tjn[st.lane_owner[x]]):] for x in range(len(tln))]
        fromj = g["lane_from_junction"]
        toj = g["lane_to_junction"]
        internal = (g["lane_flags"] & LF_INTERNAL) != 0
        lane_names = [(":" + jname[toj[l]] + suffix[lane_t[l]]) if internal[l] else (jname[fromj[l]] + jname[toj[l]] + "_0") for l in range(nL)]
        g["lane_name_begin"], g["lane_names"] = _pack_names(lane_names)
        g["edge_name_begin"], g["edge_names"] = _pack_names([s[:s.rindex("_")] for s in lane_names])
        g["junction_name_begin"], g["junction_names"] = _pack_names(jname)
        g["tl_name_begin"], g["tl_names"] = _pack_names(tl_names)
        order = np.argsort(np.array(lane_names, dtype=object), kind="stable")
        rank = np.zeros(nL, np.int32)
        rank[order] = np.arange(nL, dtype=np.int32)
        g["lane_idrank"] = rank
    g["grid_n"] = np.array([n], np.int32)
    g["grid_spacing"] = np.array([spacing], np.float64)
    return g


def _junction_pos(t, st, j):
    """centre of template junction j, estimated as the mean of the end points of its incoming normal lanes' shapes mirrored
    with the start points of its outgoing ones (exact enough to measure the grid spacing, which is a multiple of 0.01)"""
    nL = len(t["lane_length"])
    pts = []
    xy = t["shape_xy"].reshape(-1, 2)
    for l in range(nL):
        if st.internal[l]:
            continue
        if t["lane_to_junction"][l] == j:
            pts.append(xy[t["lane_shape_begin"][l + 1] - 1])
        if t["lane_from_junction"][l] == j:
            pts.append(xy[t["lane_shape_begin"][l]])
    p = np.mean(np.array(pts), axis=0)
    return np.round(p, 2)


def set_tls_offsets(g, offsets_ms, begin_ms=0):
    """gives every static program an offset like <tlLogic offset="..."> does: the initial phase and the first switch time
    (NLJunctionControlBuilder::closeTrafficLightLogic, src/netload/NLJunctionControlBuilder.cpp:238-252)"""
    pb = g["tl_phase_begin"]
    dur = g["phase_dur"]
    cur = g["tl_cur_phase"].copy()
    nxt = g["tl_next_switch"].copy()
    for i in range(len(cur)):
        d = dur[pb[i]:pb[i + 1]]
        total = int(d.sum())
        if total == 0:
            continue
        off = int(offsets_ms[i])
        if off >= 0:
            o = (begin_ms + total - (off % total)) % total
        else:
            o = (begin_ms + ((-off) % total)) % total
        step = 0
        while o >= d[step]:
            o -= int(d[step])
            step += 1
        cur[i] = step
        nxt[i] = int(d[step]) - o + begin_ms
    g["tl_cur_phase"] = cur
    g["tl_next_switch"] = nxt
    return g


def make_demand(g, n_vehicles, route_edges, depart_end_s, seed=42, sigma=None, speeddev=0.1, names=False):
    """n_vehicles vehicles on random-walk routes of `route_edges` edges without U-turns, start edges uniform over the normal
    edges, departures uniform in [0, depart_end_s) sorted (the order the reference's loader keeps), default vType.  Adds the
    route_* / veh_* arrays to g.  speeddev: the vType's speedFactor deviation, sampled like normc(1, dev, 0.2, 2) rounded to 2
    decimals (SUMOVTypeParameter.cpp:905-1040 defaults for passenger cars); 0 = every vehicle drives exactly the limit."""
    rng = np.random.default_rng(seed)
    nL = len(g["lane_length"])
    normal = np.nonzero((g["lane_flags"] & LF_INTERNAL) == 0)[0]
    # successor table without turnarounds: for every normal lane the normal lanes reachable through its links
    lb = g["lane_link_begin"].astype(np.int64)
    nxt = -np.ones((nL, 4), np.int64)
    cnt = np.zeros(nL, np.int64)
    lfrom = g["link_from"].astype(np.int64)
    from_normal = (g["lane_flags"][lfrom] & LF_INTERNAL) == 0

    def fill(mask):
        k = np.nonzero(mask)[0]
        f = lfrom[k]
        # rank of each selected link among the selected links of its lane (links of a lane are contiguous and ascending)
        first = np.ones(len(k), bool)
        first[1:] = f[1:] != f[:-1]
        start = np.maximum.accumulate(np.where(first, np.arange(len(k)), 0))
        slot = np.arange(len(k)) - start + cnt[f]
        keep = slot < 4
        nxt[f[keep], slot[keep]] = g["link_to"][k[keep]]
        np.add.at(cnt, f[keep], 1)

    fill(from_normal & (g["link_dir"] != ord("t")))
    # lanes whose only way on is the turnaround (none inside a grid; kept for other templates): allow it there
    fill(from_normal & (cnt[lfrom] == 0))
    cur = rng.choice(normal, size=n_vehicles)
    routes = np.zeros((n_vehicles, route_edges), np.int32)
    routes[:, 0] = cur
    for s in range(1, route_edges):
        pick = (rng.random(n_vehicles) * cnt[cur]).astype(np.int64)
        cur = nxt[cur, pick]
        routes[:, s] = cur
    g["route_begin"] = (np.arange(n_vehicles + 1, dtype=np.int64) * route_edges).astype(np.int32) if n_vehicles * route_edges < 2 ** 31 else None
    if g["route_begin"] is None:
        raise ValueError("route table too large for int32 offsets")
    g["route_edges"] = routes.reshape(-1)
    depart = np.sort((rng.random(n_vehicles) * depart_end_s * 1000.).astype(np.int64))
    g["veh_depart"] = depart
    g["veh_type"] = np.zeros(n_vehicles, np.int32)
    g["veh_route"] = np.arange(n_vehicles, dtype=np.int32)
    g["veh_numid"] = np.arange(n_vehicles, dtype=np.int32)
    if speeddev > 0:
        sf = rng.normal(1.0, speeddev, n_vehicles)
        bad = (sf < 0.2) | (sf > 2.0)
        while bad.any():
            sf[bad] = rng.normal(1.0, speeddev, int(bad.sum()))
            bad = (sf < 0.2) | (sf > 2.0)
        g["veh_speedfactor"] = np.round(sf, 2)
        g["opt_max_speed_factor"] = np.array([1.0])
    else:
        g["veh_speedfactor"] = np.ones(n_vehicles)
    g["veh_arrivalpos"] = g["lane_length"][routes[:, -1]].copy()
    g["veh_departpos_proc"] = np.zeros(n_vehicles, np.int32)
    g["veh_departpos"] = np.zeros(n_vehicles)
    g["veh_departspeed_proc"] = np.zeros(n_vehicles, np.int32)
    g["veh_departspeed"] = -np.ones(n_vehicles)
    g["veh_departlane_proc"] = np.zeros(n_vehicles, np.int32)
    g["veh_departlane"] = np.zeros(n_vehicles, np.int32)
    g["veh_emissions"] = np.zeros(n_vehicles, np.uint8)
    if names:
        vn = ["v%d" % i for i in range(n_vehicles)]
        g["veh_name_begin"], g["veh_names"] = _pack_names(vn)
        order = np.argsort(np.array(vn, dtype=object), kind="stable")
        rank = np.zeros(n_vehicles, np.int32)
        rank[order] = np.arange(n_vehicles, dtype=np.int32)
        g["veh_idrank"] = rank
    else:
        g["veh_idrank"] = _decimal_string_rank(n_vehicles)
    if sigma is not None:
        g["vt_sigma"] = np.full(g["vt_sigma"].shape, float(sigma))
    # slot order: by departure lane, then departure time.  The slot is only the index of the vehicle's state in HBM (identity =
    # veh_numid / veh_idrank, both kept); vehicles that start on the same lane drive near each other for a long time, so their state
    # shares sectors, and in a partitioned run the slots a rank touches are (initially) one contiguous range instead of every N-th.
    # Per departure lane the order stays the load order, which is all MSInsertionControl's semantics need (sb2_build.h dq_*).
    perm = np.argsort(routes[:, 0], kind="stable")
    for k in list(g):
        if k.startswith("veh_") and k not in ("veh_names", "veh_name_begin") and len(g[k]) == n_vehicles:
            g[k] = g[k][perm]
    if names:
        vn = [vn[i] for i in perm]
        g["veh_name_begin"], g["veh_names"] = _pack_names(vn)
    return g


def _decimal_string_rank(n):
    """rank of the ids "v0", "v1", ... "v<n-1>" in std::string order (MSVehicle::isLeader's id tie breaker, FCD output order)
    without building the strings: compare by (digits left-aligned) -> sort key = number scaled to the longest width, then length"""
    i = np.arange(n, dtype=np.int64)
    width = np.ones(n, np.int64)
    p = 10
    while p <= max(n - 1, 1):
        width += i >= p
        p *= 10
    maxw = int(width.max()) if n > 0 else 1
    key = i * (10 ** (maxw - width))
    order = np.lexsort((width, key))
    rank = np.zeros(n, np.int32)
    rank[order] = np.arange(n, dtype=np.int32)
    return rank


def write_routes_xml(g, path, vtype_attrs='sigma="0.5"'):
    """the demand of make_demand as a .rou.xml for the reference (edge ids from the generated names)"""
    en = flat_names(g, "edge")
    rb = g["route_begin"]
    re_ = g["route_edges"]
    with open(path, "w") as f:
        f.write('<routes>\n    <vType id="car" %s/>\n' % vtype_attrs)
        for v in np.argsort(g["veh_numid"], kind="stable"):      # load order = departure order (the slots are sorted by lane)
            r = g["veh_route"][v]
            f.write('    <vehicle id="v%d" type="car" depart="%.3f"><route edges="%s"/></vehicle>\n' % (
                g["veh_numid"][v], g["veh_depart"][v] / 1000., " ".join(en[e] for e in re_[rb[r]:rb[r + 1]])))
        f.write("</routes>\n")
