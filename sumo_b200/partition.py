"""Edge-partitioned multi-GPU run of ONE connected network (device side: sumo_b200/csrc/sb2_halo.h).

The junction grid is cut into strips of columns, one strip per rank (one process per GPU).  A junction with its internal lanes
and the normal lanes ending at it is a block; a strip owns whole blocks.  Every rank creates the engine from the whole scenario
and tells it, per lane, whether it is its own (3), ghost (2: the neighbour's first column, simulated redundantly), halo (1: the
neighbour's second column, read only) or invisible (0).  After every step the ranks swap the final state of their two border
columns with each neighbour: grouped NCCL send / recv through torch.distributed (plumbing) on device buffers the engine packs
and unpacks itself, plus one 3-value MAX all-reduce for the fleet aggregates behind MSLane::getMaximumBrakeDist.

plan_partition is pure numpy and is what the CPU tests exercise (world_size 2 over gloo for the transport, tests/host_emul for the
bit-exactness of the scheme); PartitionedEngine needs CUDA.
"""
import ctypes

import numpy as np

from .flatfile import names as flat_names

LV_NONE, LV_HALO, LV_GHOST, LV_AUTH = 0, 1, 2, 3


def junction_columns(g):
    """column index of every junction of a grid scenario: gridgen scenarios carry grid_n (junction = column * n + row); netgenerate
    grids are recognised by their alphanumerical junction ids (column letters + row number, src/netgen/NGNet.cpp:80-99)"""
    nj = int(max(g["lane_to_junction"].max(), g["lane_from_junction"].max())) + 1
    if "grid_n" in g:
        n = int(g["grid_n"][0])
        return np.arange(nj) // n, n
    jn = flat_names(g, "junction")
    cols = np.zeros(len(jn), np.int64)
    for j, s in enumerate(jn):
        k = 0
        while k < len(s) and s[k].isalpha() and s[k].isupper():
            k += 1
        if k == 0 or k == len(s) or not s[k:].isdigit():
            raise ValueError("junction id '%s' is not a netgenerate grid id: cannot derive the column strips" % s)
        c = 0
        for ch in s[:k]:
            c = c * 26 + ord(ch) - ord("A")
        cols[j] = c
    return cols, int(cols.max()) + 1


def column_ranges(n_columns, world):
    """contiguous, balanced split of the columns; every rank needs at least two (it exports two columns to each side)"""
    if world < 1 or n_columns < 2 * world:
        raise ValueError("%d columns cannot be cut into %d strips of at least 2 columns" % (n_columns, world))
    q, r = divmod(n_columns, world)
    out, b = [], 0
    for k in range(world):
        e = b + q + (1 if k < r else 0)
        out.append((b, e))
        b = e
    return out


def plan_partition(g, rank, world, capacity=None):
    """what rank `rank` of `world` does with every lane.  Returns dict(lane_vis, exp=[left, right], imp=[left, right], capacity,
    columns=(c0, c1))."""
    cols, ncol = junction_columns(g)
    c0, c1 = column_ranges(ncol, world)[rank]
    lane_col = cols[g["lane_to_junction"]]          # block of a lane = the junction it ends at (internal lanes: their junction)
    vis = np.zeros(len(lane_col), np.uint8)
    vis[(lane_col >= c0) & (lane_col < c1)] = LV_AUTH
    vis[(lane_col == c0 - 1) | (lane_col == c1)] = LV_GHOST
    vis[(lane_col == c0 - 2) | (lane_col == c1 + 1)] = LV_HALO
    lanes = np.arange(len(lane_col), dtype=np.int32)
    empty = np.zeros(0, np.int32)
    exp = [lanes[(lane_col >= c0) & (lane_col < c0 + 2)] if rank > 0 else empty,
           lanes[(lane_col >= c1 - 2) & (lane_col < c1)] if rank < world - 1 else empty]
    imp = [lanes[(lane_col >= c0 - 2) & (lane_col < c0)] if rank > 0 else empty,
           lanes[(lane_col >= c1) & (lane_col < c1 + 2)] if rank < world - 1 else empty]
    if capacity is None:
        # vehicles of two columns at three times the average density, identical on every rank
        capacity = int(3 * 2 * len(g["veh_type"]) / max(ncol, 1)) + 4096
    return {"lane_vis": vis, "exp": exp, "imp": imp, "capacity": int(capacity), "columns": (c0, c1), "n_columns": ncol}


def write_partition_file(path, g, world, capacity=None):
    """the plans of all ranks for tests/host_emul --partitions (int32 words, layout in emul_main.cpp)"""
    plans = [plan_partition(g, r, world, capacity) for r in range(world)]
    words = [world, len(plans[0]["lane_vis"]), plans[0]["capacity"]]
    for p in plans:
        words += p["lane_vis"].astype(np.int32).tolist()
        for side in range(2):
            words += [len(p["exp"][side])] + p["exp"][side].tolist() + [len(p["imp"][side])] + p["imp"][side].tolist()
    np.array(words, np.int32).tofile(path)
    return plans


def exchange_messages(dist, rank, world, send, recv):
    """one grouped point-to-point exchange with both neighbours: send[side] goes to rank - 1 (side 0) / rank + 1 (side 1), recv[side]
    is filled by that neighbour.  Tensors may live on the GPU (NCCL) or on the host (gloo: the CPU tests)."""
    ops = []
    for side, peer in ((0, rank - 1), (1, rank + 1)):
        if 0 <= peer < world:
            ops.append(dist.P2POp(dist.isend, send[side], peer))
            ops.append(dist.P2POp(dist.irecv, recv[side], peer))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()


class _DevPtr:
    """wraps a raw device pointer for torch.as_tensor (CUDA array interface)"""

    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 3}


class PartitionedEngine:
    """one rank's engine of a partitioned run; same calls as engine.Engine where bench.py needs them"""

    def __init__(self, scenario, dist, rank, world, device=0, capacity=None):
        import torch
        from .engine import Engine
        self.torch, self.dist, self.rank, self.world = torch, dist, rank, world
        self.plan = plan_partition(scenario, rank, world, capacity)
        self.eng = Engine(scenario, device=device)
        lib, h = self.eng._lib, self.eng._h
        lib.sumob200_set_partition.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p), ctypes.c_void_p,
                                               ctypes.POINTER(ctypes.c_void_p), ctypes.c_void_p, ctypes.c_int]
        lib.sumob200_halo_message_bytes.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
        lib.sumob200_halo_message_bytes.restype = ctypes.c_int64
        lib.sumob200_halo_pack.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
        lib.sumob200_halo_unpack.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
        lib.sumob200_stream.argtypes = [ctypes.c_void_p]
        lib.sumob200_stream.restype = ctypes.c_void_p
        lib.sumob200_fleet_aggregates.argtypes = [ctypes.c_void_p]
        lib.sumob200_fleet_aggregates.restype = ctypes.c_void_p
        vis = np.ascontiguousarray(self.plan["lane_vis"])
        self._keep = [np.ascontiguousarray(a, dtype=np.int32) for a in self.plan["exp"] + self.plan["imp"]]
        exp_p = (ctypes.c_void_p * 2)(self._keep[0].ctypes.data, self._keep[1].ctypes.data)
        imp_p = (ctypes.c_void_p * 2)(self._keep[2].ctypes.data, self._keep[3].ctypes.data)
        n_exp = np.array([len(self._keep[0]), len(self._keep[1])], np.int32)
        n_imp = np.array([len(self._keep[2]), len(self._keep[3])], np.int32)
        from .engine import _check
        _check(lib, lib.sumob200_set_partition(h, vis.ctypes.data, exp_p, n_exp.ctypes.data, imp_p, n_imp.ctypes.data, self.plan["capacity"]), "sumob200_set_partition")
        self.dev = torch.device("cuda", device)
        self.stream = torch.cuda.ExternalStream(lib.sumob200_stream(h), device=self.dev)
        self.send, self.recv = [None, None], [None, None]
        for side, peer in ((0, rank - 1), (1, rank + 1)):
            if 0 <= peer < world:
                self.send[side] = torch.zeros(lib.sumob200_halo_message_bytes(h, side, 1), dtype=torch.uint8, device=self.dev)
                self.recv[side] = torch.zeros(lib.sumob200_halo_message_bytes(h, side, 0), dtype=torch.uint8, device=self.dev)
        self.fleet = torch.as_tensor(_DevPtr(lib.sumob200_fleet_aggregates(h), 3, "<f8"), device=self.dev)
        self.num_vehicles = self.eng.num_vehicles
        self.num_lanes = self.eng.num_lanes
        self.exchanges = 0
        self.bytes_sent = 0
        import os as _os
        self._prof = [] if _os.environ.get("SUMOB200_HALO_PROF") else None
        self._prof_ms, self._prof_n = {}, 0
        self._skip_fleet = bool(_os.environ.get("SUMOB200_NO_FLEET_SYNC"))   # measurement only: results are then not exact

    # ---- the step: engine step + halo exchange, all enqueued on the engine's stream
    HEADER, LANE_REC, VEH_REC = 16, 64, 288    # sizeof HaloHeader / HaloLane / HaloVeh (static_assert in sb2_halo.h)

    def _limit(self, last_count):
        """vehicle records both ends transport in the next message, from the count of the previous one (which both know: the
        sender packed it, the receiver read it in the header): 25 % + 512 head room, at most the capacity"""
        if last_count is None:
            return self.plan["capacity"]
        return int(min(self.plan["capacity"], last_count + last_count // 4 + 512))

    def _mark(self, name):
        """SUMOB200_HALO_PROF=1: CUDA events between the parts of the exchange (summed in halo_stats()['ms_per_step'])"""
        if self._prof is None:
            return
        ev = self.torch.cuda.Event(enable_timing=True)
        ev.record(self.stream)
        self._prof.append((name, ev))

    def _exchange(self):
        from .engine import _check
        lib, h, torch = self.eng._lib, self.eng._h, self.torch
        if not hasattr(self, "_counts"):
            # per side: [pinned host copy of the last sent header, of the last received header, event]
            self._counts = [{"sent": torch.zeros(4, dtype=torch.int32).pin_memory(), "recv": torch.zeros(4, dtype=torch.int32).pin_memory(),
                             "event": None} for _ in (0, 1)]
        with torch.cuda.stream(self.stream):
            self._mark("step")
            send_v, recv_v = [None, None], [None, None]
            for side in (0, 1):
                if self.send[side] is None:
                    continue
                st = self._counts[side]
                if st["event"] is not None:
                    st["event"].synchronize()            # recorded a whole step ago
                    lim_s, lim_r = self._limit(int(st["sent"][1])), self._limit(int(st["recv"][1]))
                else:
                    lim_s = lim_r = self.plan["capacity"]
                n_exp, n_imp = len(self._keep[side]), len(self._keep[2 + side])
                _check(lib, lib.sumob200_halo_pack(h, side, self.send[side].data_ptr(), lim_s), "sumob200_halo_pack")
                send_v[side] = self.send[side][:self.HEADER + self.LANE_REC * n_exp + self.VEH_REC * lim_s]
                recv_v[side] = self.recv[side][:self.HEADER + self.LANE_REC * n_imp + self.VEH_REC * lim_r]
                self.bytes_sent += send_v[side].numel()
            self._mark("pack")
            exchange_messages(self.dist, self.rank, self.world, send_v, recv_v)
            self._mark("transfer")
            for side in (0, 1):
                if self.recv[side] is not None:
                    _check(lib, lib.sumob200_halo_unpack(h, side, self.recv[side].data_ptr()), "sumob200_halo_unpack")
                    st = self._counts[side]
                    st["sent"].copy_(self.send[side][:16].view(torch.int32), non_blocking=True)
                    st["recv"].copy_(self.recv[side][:16].view(torch.int32), non_blocking=True)
                    st["event"] = torch.cuda.Event()
                    st["event"].record(self.stream)
            # fleet aggregates: max speedFactor, MIN deceleration, max minGap over the departed vehicles of the whole network
            self._mark("unpack")
            if not self._skip_fleet:
                self.fleet[1].neg_()
                self.dist.all_reduce(self.fleet, op=self.dist.ReduceOp.MAX)
                self.fleet[1].neg_()
            self._mark("allreduce")
        self.exchanges += 1
        if self._prof is not None and len(self._prof) >= 5 * 64:
            self._fold_prof()

    def _fold_prof(self):
        self.stream.synchronize()
        prev = None
        for name, ev in self._prof:
            if prev is not None and name != "step":
                self._prof_ms[name] = self._prof_ms.get(name, 0.) + prev.elapsed_time(ev)
            elif prev is not None:
                self._prof_ms["step"] = self._prof_ms.get("step", 0.) + prev.elapsed_time(ev)
            prev = ev
        self._prof_n += sum(1 for name, _ in self._prof if name == "step")
        self._prof = []

    def step(self, n=1, result=True):
        for _ in range(n):
            self.eng.step(1, result=False)
            self._exchange()
        return self.eng.step(0) if result else None

    def step_timed(self, n):
        torch = self.torch
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(self.stream):
            self.stream.synchronize()
            a.record(self.stream)
            self.step(n, result=False)
            b.record(self.stream)
            b.synchronize()
        return self.eng.step(0), a.elapsed_time(b)

    def vehicles(self):
        """state of the vehicles this rank owns (lane < 0 for everything else)"""
        st = self.eng.vehicles().copy()
        lane = st["lane"]
        own = np.zeros(len(st), bool)
        on = lane >= 0
        own[on] = self.plan["lane_vis"][lane[on]] == LV_AUTH
        st["lane"][on & ~own] = -5
        return st

    def halo_stats(self):
        torch = self.torch
        self.stream.synchronize()
        out = {"exchanges": self.exchanges, "columns": list(self.plan["columns"]), "capacity_records": self.plan["capacity"],
               "bytes_sent_per_step": self.bytes_sent // max(self.exchanges, 1)}
        if self._prof is not None:
            self._fold_prof()
            out["ms_per_step"] = {k: v / max(self._prof_n, 1) for k, v in self._prof_ms.items()}
        for side, name in ((0, "left"), (1, "right")):
            if self.send[side] is not None:
                hdr = self.send[side][:16].cpu().numpy().view(np.int32)
                out[name] = {"lanes": int(hdr[0]), "vehicle_records": int(hdr[1]), "message_bytes": int(self.send[side].numel()),
                             "used_bytes": int(16 + 64 * hdr[0] + 288 * hdr[1])}   # sizeof HaloHeader / HaloLane / HaloVeh (sb2_halo.h)
        return out

    # ---- pass-throughs
    def fcd_begin(self):
        self.eng.fcd_begin()

    def fcd_end(self):
        return self.eng.fcd_end()

    def set_profiling(self, enabled):
        self.eng.set_profiling(enabled)

    def phase_times(self, reset=True):
        return self.eng.phase_times(reset)

    def _merge_owned(self, rows, det_lane):
        """detector records of all ranks merged: a detector is reported by the rank that owns its lane, the rows of the others are
        zeroed and the byte-wise sum over the ranks (one all-reduce) is the owner's row"""
        torch = self.torch
        own = self.plan["lane_vis"][np.asarray(det_lane)] == LV_AUTH
        raw = np.ascontiguousarray(rows).view(np.uint8).reshape(len(rows), rows.dtype.itemsize).copy()
        raw[~own] = 0
        buf = torch.from_numpy(raw).to(self.dev)
        self.dist.all_reduce(buf)
        return buf.cpu().numpy().reshape(-1).view(rows.dtype).copy()

    def e1_interval(self, begin_ms, end_ms, det_lane):
        """Engine.e1_interval for the whole network (det_lane = the scenario's e1_lane array); collective: every rank calls it"""
        return self._merge_owned(self.eng.e1_interval(begin_ms, end_ms), det_lane)

    def e2_interval(self, det_lane):
        """Engine.e2_interval for the whole network (det_lane = the scenario's e2_lane array); collective: every rank calls it"""
        return self._merge_owned(self.eng.e2_interval(), det_lane)

    @property
    def kernel_launches(self):
        return self.eng.kernel_launches

    @property
    def time_ms(self):
        return self.eng.time_ms

    def close(self):
        # torch buffers that were used on the engine's stream must go before the stream does (their deleter records on it)
        torch = self.torch
        self.stream.synchronize()
        self.send, self.recv, self.fleet = [None, None], [None, None], None
        if hasattr(self, "_counts"):
            del self._counts
        torch.cuda.synchronize()
        torch.cuda.empty_cache()
        self.stream = None
        self.eng.close()


class LocalPartitionedRun:
    """all strips of a partitioned run as engines of ONE process on ONE GPU, messages handed over by device pointer: the device
    side of the scheme (sumob200_set_partition, halo pack / unpack kernels) without NCCL -- used by the single-GPU test suite.
    (Separate processes sharing one GPU must not wait on each other, B200_PROFILING.md; here the engines are stepped in turn.)"""

    def __init__(self, scenario, world, device=0, capacity=None):
        import torch
        from .engine import Engine, _check
        self.torch, self.world = torch, world
        self.parts = []
        dev = torch.device("cuda", device)
        for rank in range(world):
            plan = plan_partition(scenario, rank, world, capacity)
            eng = Engine(scenario, device=device)
            lib, h = eng._lib, eng._h
            lib.sumob200_set_partition.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p), ctypes.c_void_p,
                                                   ctypes.POINTER(ctypes.c_void_p), ctypes.c_void_p, ctypes.c_int]
            lib.sumob200_halo_message_bytes.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
            lib.sumob200_halo_message_bytes.restype = ctypes.c_int64
            lib.sumob200_halo_pack.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
            lib.sumob200_halo_unpack.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
            lib.sumob200_fleet_aggregates.argtypes = [ctypes.c_void_p]
            lib.sumob200_fleet_aggregates.restype = ctypes.c_void_p
            keep = [np.ascontiguousarray(a, dtype=np.int32) for a in plan["exp"] + plan["imp"]]
            vis = np.ascontiguousarray(plan["lane_vis"])
            exp_p = (ctypes.c_void_p * 2)(keep[0].ctypes.data, keep[1].ctypes.data)
            imp_p = (ctypes.c_void_p * 2)(keep[2].ctypes.data, keep[3].ctypes.data)
            n_exp = np.array([len(keep[0]), len(keep[1])], np.int32)
            n_imp = np.array([len(keep[2]), len(keep[3])], np.int32)
            _check(lib, lib.sumob200_set_partition(h, vis.ctypes.data, exp_p, n_exp.ctypes.data, imp_p, n_imp.ctypes.data, plan["capacity"]), "sumob200_set_partition")
            send = [None, None]
            for side, peer in ((0, rank - 1), (1, rank + 1)):
                if 0 <= peer < world:
                    send[side] = torch.zeros(lib.sumob200_halo_message_bytes(h, side, 1), dtype=torch.uint8, device=dev)
            fleet = torch.as_tensor(_DevPtr(lib.sumob200_fleet_aggregates(h), 3, "<f8"), device=dev)
            self.parts.append({"eng": eng, "plan": plan, "send": send, "fleet": fleet})

    def step(self):
        from .engine import _check
        torch = self.torch
        for p in self.parts:
            p["eng"].step(1, result=False)
        for p in self.parts:
            lib, h = p["eng"]._lib, p["eng"]._h
            for side in (0, 1):
                if p["send"][side] is not None:
                    _check(lib, lib.sumob200_halo_pack(h, side, p["send"][side].data_ptr(), 0), "sumob200_halo_pack")
        torch.cuda.synchronize()
        for rank, p in enumerate(self.parts):
            lib, h = p["eng"]._lib, p["eng"]._h
            for side, peer in ((0, rank - 1), (1, rank + 1)):
                if 0 <= peer < self.world:
                    _check(lib, lib.sumob200_halo_unpack(h, side, self.parts[peer]["send"][side ^ 1].data_ptr()), "sumob200_halo_unpack")
        torch.cuda.synchronize()
        agg = torch.stack([p["fleet"] for p in self.parts])
        red = torch.stack([agg[:, 0].max(), agg[:, 1].min(), agg[:, 2].max()])
        for p in self.parts:
            p["fleet"].copy_(red)
        torch.cuda.synchronize()

    def results(self):
        return [p["eng"].step(0) for p in self.parts]

    def vehicles(self):
        """merged state: every vehicle as its owner reports it (lane -5 where no strip owns it: not on the road)"""
        out = None
        for p in self.parts:
            st = p["eng"].vehicles().copy()
            lane = st["lane"]
            on = lane >= 0
            own = np.zeros(len(st), bool)
            own[on] = p["plan"]["lane_vis"][lane[on]] == LV_AUTH
            if out is None:
                out = st.copy()
                out["lane"][:] = -5
            assert not (own & (out["lane"] >= 0)).any(), "a vehicle is owned by two strips"
            out[own] = st[own]
        return out

    def close(self):
        for p in self.parts:
            p["eng"].close()
