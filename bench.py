#!/usr/bin/env python3
"""Throughput benchmark of the step engine: vehicle-updates/s (SUMO's UPS numerator, MSNet.cpp:959).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--grid G] [--vehicles V] [--sigma S]     our arm (libsumob200.so, C ABI)
  python bench.py --impl reference ...                                                        the reference's own CPU path

Workload = BASELINE.json configs[1] as SURVEY.md section 8d defines it (C2): ONE connected `netgenerate --grid --grid.number=100
--grid.length=800 --default-junction-type traffic_light` network (39,600 single-lane edges, 31,680 lane-km), 1,000,000 vehicles
(32 veh/lane-km) with Krauss sigma 0.5 and the default speedFactor deviation on random-walk routes of 40 edges, departures spread
over the first 300 s, every traffic light with its own offset (so the cost of a step is not periodic), the reference's DEFAULT
options (time-to-teleport 300 s, collision.action teleport).  The network is generated directly as arrays (sumo_b200/gridgen.py,
validated against netgenerate + the reference loader in tests/test_gridgen.py), the demand likewise (no per-vehicle host objects).
A "step" is one simulation step (DELTA_T = 1 s).

N > 1 (torchrun): ONE connected grid cut into N column strips, one per GPU, with a halo exchange over NCCL every step
(sumo_b200/partition.py, csrc/sb2_halo.h).  Default = weak scaling: the grid grows with N at the same density and per-GPU work
(side 100 * sqrt(N): 141 / 200 / 283 junctions, 2 / 4 / 8 M vehicles; N = 8 is close to BASELINE.json configs[3]); `--strong` keeps
the 100 x 100 / 1 M network for every N; `--replicas` runs N unconnected copies (no exchange).
"""
import argparse
import json
import os
import re
import shutil
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

TEMPLATE = os.path.join(ROOT, "tests", "golden", "grid5_tls800.flat")
GRID = 100
VEHICLES = 1000000
ROUTE_EDGES = 40
DEPART_END_S = 300
SPINUP_STEPS = 420          # all departures are due within 300 s; the backlog of the busiest lanes is gone soon after
PROF_STEPS = 60             # window of the per-kernel (roofline) timing
TLS_CYCLE_S = 90
REF_GRID = 12               # the reference arm simulates a 12 x 12 cut of the same workload per process (same density, same options)
# --config c4: BASELINE.json configs[3] as SURVEY.md section 8d defines it (C4): 1000 x 1000 grid at 100 m, priority junctions,
# 10 M vehicles (25 veh/lane-km), Krauss sigma 0; fixed size for every N (edge-partitioned across the GPUs: strong scaling)
WORKLOADS = {
    "c2": {"template": "grid5_tls800", "grid": GRID, "vehicles": VEHICLES, "route_edges": 40, "depart_end": 300, "spinup": 420, "length": 800,
           "jtype": "traffic_light", "ref_grid": 12, "sigma": 0.5, "tls": True},
    "c4": {"template": "grid5_prio100", "grid": 1000, "vehicles": 10000000, "route_edges": 48, "depart_end": 100, "spinup": 200, "length": 100,
           "jtype": "priority", "ref_grid": 40, "sigma": 0.0, "tls": False},
}


def use_workload(name):
    """selects the workload constants (module globals: the functions below read them)"""
    global TEMPLATE, ROUTE_EDGES, DEPART_END_S, SPINUP_STEPS, REF_GRID, GRID_LENGTH, JUNCTION_TYPE, HAS_TLS
    w = WORKLOADS[name]
    TEMPLATE = os.path.join(ROOT, "tests", "golden", w["template"] + ".flat")
    ROUTE_EDGES, DEPART_END_S, SPINUP_STEPS, REF_GRID = w["route_edges"], w["depart_end"], w["spinup"], w["ref_grid"]
    GRID_LENGTH, JUNCTION_TYPE, HAS_TLS = w["length"], w["jtype"], w["tls"]
    return w


GRID_LENGTH, JUNCTION_TYPE, HAS_TLS = 800, "traffic_light", True
REF_DENSITY_GRID, REF_DENSITY_VEHICLES = GRID, VEHICLES
# algorithmic bytes per vehicle-update (SURVEY.md section 8d / DESIGN.md "roofline"): whole step, and the plan kernel alone
BYTES_PER_UPDATE_STEP = 576
BYTES_PER_UPDATE_PLAN = 320
TRAFFIC_FILE = os.path.join(ROOT, "profiles", "r02_k_plan_traffic.json")   # dram__bytes of one k_plan launch from an ncu --set full capture of THIS kernel


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def tls_offsets(n_tl, seed=7):
    return np.random.default_rng(seed).integers(0, TLS_CYCLE_S, n_tl) * 1000


def build_scenario(grid, vehicles, sigma, emissions=False, names=False, seed=1):
    from sumo_b200 import gridgen
    from sumo_b200.flatfile import load_flat
    g = gridgen.make_grid(load_flat(TEMPLATE), grid, with_names=grid <= 200)
    if HAS_TLS:
        gridgen.set_tls_offsets(g, tls_offsets(len(g["tl_cur_phase"])))
    gridgen.make_demand(g, vehicles, ROUTE_EDGES, DEPART_END_S, seed=seed, sigma=sigma, speeddev=0.1, names=names)
    if emissions:
        g["veh_emissions"] = np.ones(vehicles, np.uint8)
    return g


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region (B200_PROFILING.md's clocks line).  NVML is polled from a
    thread every 5 ms (the timed region of the default run is ~0.1 s, shorter than nvidia-smi's start-up); nvidia-smi -lms is
    the fallback when the NVML binding is missing."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    REASONS = ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"), (0x4, "sw_power_cap"))

    def __init__(self, index):
        self.index = index
        self.sm, self.smax, self.reasons = [], [], set()
        self.proc = None
        self.nvml = None
        self.handle = None
        self.running = False
        self.thread = None
        self.source = None
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = index
            if vis:
                ids = [x.strip() for x in vis.split(",") if x.strip()]
                if index < len(ids) and ids[index].isdigit():
                    phys = int(ids[index])
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.smax_static = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
            self.source = "nvml"
        except Exception:
            self.nvml = None

    def _sample_nvml(self):
        n = self.nvml
        try:
            self.sm.append(float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)))
            self.smax.append(self.smax_static)
            try:
                bits = n.nvmlDeviceGetCurrentClocksEventReasons(self.handle)
            except Exception:
                bits = n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
            for bit, name in self.REASONS:
                if bits & bit:
                    self.reasons.add(name)
        except Exception:
            pass

    def _poll(self):
        while self.running:
            self._sample_nvml()
            time.sleep(0.005)

    def _read(self):
        for line in self.proc.stdout:
            f = [x.strip() for x in line.strip().split(",")]
            if len(f) < 7:
                continue
            try:
                self.sm.append(float(f[0]))
                self.smax.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    self.reasons.add(name)

    def start(self):
        if self.nvml is not None:
            self.running = True
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
            self.source = "nvidia-smi"
        except Exception:
            self.proc = None

    def stop(self):
        if self.nvml is not None:
            self._sample_nvml()           # at least one sample taken right at the end of the timed region
            self.running = False
            self.thread.join()
        elif self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()
        else:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"], "samples": 0}
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": max(self.smax) if self.smax else None,
                "reasons": sorted(self.reasons), "samples": len(self.sm), "source": self.source}


def reference_ups(total_steps, sigma, procs=None, grid=None):
    """the reference's own CPU implementation (oracle/_ref/sumo: the unmodified sources, oracle/build_ref.sh) on a `grid` x `grid`
    cut of the bench workload: the same netgenerate options, density (32 veh/lane-km), route length, departure window, vType,
    per-junction signal offsets and default sumo options; the net comes from oracle/_ref/netgenerate, the demand from the same
    generator as a .rou.xml.  `sumo --threads` needs the FOX toolkit (src/microsim/MSFrame.cpp:1119-1124), which is absent, so one
    sumo process is single-threaded; all host cores are used by running `procs` processes side by side (each the whole cut).
    Aggregate = sum of the processes' own UPS (vehicle updates / their simulation wall time, MSNet.cpp:959)."""
    from sumo_b200 import gridgen
    from sumo_b200.flatfile import load_flat, names
    exe = os.path.join(ROOT, "oracle", "_ref", "sumo")
    netgen = os.path.join(ROOT, "oracle", "_ref", "netgenerate")
    if not (os.path.exists(exe) and os.path.exists(netgen)):
        return None
    procs = procs or max(1, os.cpu_count() or 1)
    grid = grid or REF_GRID
    tmp = tempfile.mkdtemp(prefix="sumob200_ref_")
    try:
        net = os.path.join(tmp, "ref.net.xml")
        subprocess.check_call([netgen, "--grid", "--grid.number=%d" % grid, "--grid.length=%d" % GRID_LENGTH, "--default-junction-type", JUNCTION_TYPE,
                               "-o", net, "--no-warnings"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        g = gridgen.make_grid(load_flat(TEMPLATE), grid)
        if HAS_TLS:
            off = dict(zip(names(g, "tl"), tls_offsets(len(g["tl_cur_phase"])) // 1000))
            text = re.sub(r'<tlLogic id="([^"]+)" type="static" programID="0" offset="0">',
                          lambda m: '<tlLogic id="%s" type="static" programID="0" offset="%d">' % (m.group(1), off[m.group(1)]), open(net).read())
            open(net, "w").write(text)
        vehicles = int(round(REF_DENSITY_VEHICLES * (grid * (grid - 1)) / float(REF_DENSITY_GRID * (REF_DENSITY_GRID - 1))))
        gridgen.make_demand(g, vehicles, ROUTE_EDGES, DEPART_END_S, seed=1, sigma=sigma, speeddev=0.1)
        rou = os.path.join(tmp, "ref.rou.xml")
        # speedFactor: the reference samples its own normc(1, 0.1, 0.2, 2) values from the vType distribution
        gridgen.write_routes_xml(g, rou, 'sigma="%g"' % sigma)
        cmd = [exe, "-n", net, "-r", rou, "--no-step-log", "--no-warnings", "--default.departspeed", "0", "--default.departlane", "first",
               "--end", str(total_steps), "--duration-log.statistics"]
        t0 = time.time()
        ps = [subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for _ in range(procs)]
        total, dur = 0.0, []
        for p in ps:
            out = p.communicate()[0]
            m = re.search(r"UPS: ([0-9.]+)", out)
            d = re.search(r"Duration: ([0-9.]+)s", out)
            if not m:
                return None
            total += float(m.group(1))
            dur.append(float(d.group(1)) if d else 0.0)
        return {"ups": total, "ups_per_process": total / procs, "duration_s": float(np.mean(dur)), "wall_s": time.time() - t0,
                "steps": total_steps, "procs": procs, "grid": grid, "vehicles": vehicles}
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def reference_sample_text(r):
    return ("oracle/_ref/sumo (unmodified reference; single-threaded: --threads needs the FOX toolkit, absent): %d concurrent processes, one per host "
            "core, each a %dx%d cut of the workload (%d vehicles, same density / options) x %d steps from t=0, %.1f s wall; %.0f veh*steps/s per process"
            % (r["procs"], r["grid"], r["grid"], r["vehicles"], r["steps"], r["wall_s"], r["ups_per_process"]))


def workload_text(grid, vehicles, sigma, emissions):
    return ("netgenerate %dx%d grid, %d m, %s, %d vehicles on ONE connected network, Krauss sigma=%.1f, "
            "speedFactor normc(1,0.1), random-walk routes of %d edges, step-length 1 s, default teleport/collision options%s"
            % (grid, grid, GRID_LENGTH, "static traffic lights with per-junction offsets" if HAS_TLS else "priority junctions", vehicles, sigma, ROUTE_EDGES,
               ", HBEFA4 emission device on every vehicle" if emissions else ""))


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    total = SPINUP_STEPS + 180
    if args.gpus > 1 and not args.replicas and not args.strong:
        g1, v1 = args.grid, args.vehicles      # the workload our arm runs at this N (weak scaling, see main)
        args.grid = int(round(g1 * args.gpus ** 0.5))
        args.vehicles = int(round(v1 * (args.grid * (args.grid - 1)) / float(g1 * (g1 - 1))))
    r = reference_ups(total, args.sigma)
    if r is None:
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/sumo or oracle/_ref/netgenerate not built"}))
        return 0
    sample = reference_sample_text(r)
    line = {
        "impl": "reference", "metric": "vehicle-updates/s", "value": r["ups"], "unit": "veh*steps/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 * r["duration_s"] / total, "higher_is_better": True,
        "scaling": "strong" if args.gpus > 1 and args.strong and not args.replicas else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_text(args.grid, args.vehicles, args.sigma, args.emissions),
                   "reference_sample": "%dx%d cut per process (the full size needs ~2 s per step on one core)" % (r["grid"], r["grid"]),
                   "vehicles": r["vehicles"], "processes": r["procs"]},
        "cpu_baseline": {"value": r["ups"], "unit": "veh*steps/s", "cores": r["procs"], "kind": "reference", "sample": sample},
        "e2e": {"value": r["ups"], "unit": "veh*steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=60)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="c2", choices=sorted(WORKLOADS), help="c2 = BASELINE configs[1] (default), c4 = configs[3] (1000x1000 / 100 m / 10 M vehicles)")
    ap.add_argument("--grid", type=int, default=None, help="junctions per side of the connected grid")
    ap.add_argument("--vehicles", type=int, default=None, help="vehicles of the whole network")
    ap.add_argument("--sigma", type=float, default=None)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--replicas", action="store_true", help="N > 1: every rank simulates its own copy of the network instead of a strip of one network")
    ap.add_argument("--strong", action="store_true", help="N > 1: keep the N = 1 network (strong scaling) instead of growing it with N")
    ap.add_argument("--emissions", action="store_true", help="every vehicle carries the HBEFA emission device (BASELINE.json configs[4] flavour)")
    ap.add_argument("--reslot", type=int, default=64, help="single GPU: renumber the on-road vehicles in lane order every this many steps (0 = never)")
    ap.add_argument("--xml", action="store_true", help="also time the e2e loop with the FCD rows formatted as XML on the host threads")
    args = ap.parse_args()
    global REF_DENSITY_GRID, REF_DENSITY_VEHICLES
    w = use_workload(args.config)
    args.grid = args.grid or w["grid"]
    args.vehicles = args.vehicles or w["vehicles"]
    args.sigma = w["sigma"] if args.sigma is None else args.sigma
    REF_DENSITY_GRID, REF_DENSITY_VEHICLES = args.grid, args.vehicles
    if args.config == "c4":
        args.strong = True      # configs[3] is one fixed network partitioned across the GPUs
    if args.impl == "reference":
        return run_reference(args)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    from sumo_b200.engine import Engine, describe_flags

    partitioned = world > 1 and not args.replicas
    if partitioned and not args.strong:
        # weak scaling: same density, same work per GPU
        g1, v1 = args.grid, args.vehicles
        args.grid = int(round(g1 * world ** 0.5))
        args.vehicles = int(round(v1 * (args.grid * (args.grid - 1)) / float(g1 * (g1 - 1))))
    scenario = build_scenario(args.grid, args.vehicles, args.sigma, emissions=args.emissions, names=args.xml)
    if partitioned:
        from sumo_b200.partition import PartitionedEngine
        eng = PartitionedEngine(scenario, dist, rank, world, device=local_rank)
    else:
        eng = Engine(scenario, device=local_rank)
        eng.set_reslot_period(args.reslot)
    del scenario
    nV = eng.num_vehicles
    eng.step(SPINUP_STEPS, result=False)
    r0 = eng.step(args.warmup)
    assert r0.running > 0

    def barrier():
        if dist is not None:
            dist.barrier()

    # ---- device-timed region: K steps, inputs resident in HBM
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = eng.kernel_launches
    r1, ms = eng.step_timed(args.steps)
    launches = eng.kernel_launches - launches0
    clocks = sampler.stop()
    updates = r1.vehicle_updates - r0.vehicle_updates
    # ---- end-to-end through the public API with host buffers: every step's FCD rows (slot, lane, pos, speed, x, y, angle of
    # every on-road vehicle = what MSFCDExport writes) land in pinned host memory; the D2H of step i overlaps step i+1
    eng.fcd_begin()          # untimed: the engine allocates its pinned FCD buffers on first use
    eng.fcd_end()
    barrier()
    e2e_steps = args.steps
    d2h = 0
    checksum = 0.0
    t0 = time.perf_counter()
    eng.step(1, result=False)
    eng.fcd_begin()
    for _ in range(e2e_steps - 1):
        eng.step(1, result=False)
        rows = eng.fcd_end()
        d2h += rows.nbytes + 4
        checksum += float(rows["speed"][-1]) if len(rows) else 0.0
        eng.fcd_begin()
    rows = eng.fcd_end()
    d2h += rows.nbytes + 4
    rr = eng.step(0)
    e2e_s = time.perf_counter() - t0
    e2e_updates = rr.vehicle_updates - r1.vehicle_updates
    d2h = d2h // e2e_steps
    # ---- second e2e figure: the same loop with the rows formatted as the <timestep> XML of --fcd-output on the host threads
    xml = None
    if args.xml and not partitioned:
        xml_steps = max(3, min(10, args.steps))
        nbytes = 0
        t0 = time.perf_counter()
        eng.step(1, result=False)
        eng.fcd_begin()
        for i in range(xml_steps):
            if i + 1 < xml_steps:
                eng.step(1, result=False)
            rows = eng.fcd_end()
            nbytes += len(eng.fcd_xml(rows, eng.time_ms))
            if i + 1 < xml_steps:
                eng.fcd_begin()
        rx = eng.step(0)
        xs = time.perf_counter() - t0
        xml = {"value": (rx.vehicle_updates - rr.vehicle_updates) / xs, "unit": "veh*steps/s", "xml_bytes_per_step": nbytes // xml_steps, "steps": xml_steps,
               "note": "as e2e, plus sumob200_fcd_format_xml of every step's rows (byte-identical to MSFCDExport's <timestep> block) on the host threads"}
        rr = rx
    # ---- roofline of the dominant kernel (k_plan): per-phase CUDA events on the engine's stream, over a window of its own
    eng.set_profiling(True)
    eng.phase_times(reset=True)
    r2, prof_ms = eng.step_timed(PROF_STEPS)
    phases = eng.phase_times(reset=True)
    eng.set_profiling(False)
    prof_updates = r2.vehicle_updates - rr.vehicle_updates
    flags = r2.flags
    state = eng.vehicles()
    on = state["lane"] >= 0
    mean_speed = float(state["speed"][on].mean()) if on.any() else 0.0
    halo = eng.halo_stats() if partitioned else None

    eng.close()
    if dist is not None:
        import torch
        from sumo_b200.shard import reduce_step_metrics
        dev = torch.device("cuda", local_rank)
        if partitioned:
            # every rank reports the counters of the vehicles it owns: the updates add up, the time is the slowest rank's
            ms, updates = reduce_step_metrics(dist, dev, ms, updates)
            e2e_s, e2e_updates = reduce_step_metrics(dist, dev, e2e_s, e2e_updates)
        else:
            ms, updates = reduce_step_metrics(dist, dev, ms, updates)
            e2e_s, e2e_updates = reduce_step_metrics(dist, dev, e2e_s, e2e_updates)
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return 0

    peak, peak_src = measured_peak_gbs()
    value = updates / (ms / 1000.0)
    plan_ms = phases["plan"] / PROF_STEPS
    plan_gbs = BYTES_PER_UPDATE_PLAN * (prof_updates / PROF_STEPS) / (plan_ms / 1000.0) / 1e9
    traffic, traffic_src = None, None
    try:
        tr = json.load(open(TRAFFIC_FILE))
        if int(tr["vehicles_per_launch"]) == int(round(prof_updates / PROF_STEPS / 1000.0)) * 1000 or abs(tr["vehicles_per_launch"] - prof_updates / PROF_STEPS) < 0.02 * tr["vehicles_per_launch"]:
            traffic = int(tr["dram_bytes_read"] + tr["dram_bytes_write"])
            traffic_src = tr["source"]
    except Exception:
        pass
    scaling = "strong" if (partitioned and args.strong) else "weak"
    line = {
        "metric": "vehicle-updates/s", "value": value, "unit": "veh*steps/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_text(args.grid, args.vehicles, args.sigma, args.emissions),
                   "vehicles": args.vehicles, "resident": int(updates // args.steps), "lanes": int(eng.num_lanes), "mean_speed_mps": mean_speed,
                   "multi_gpu": ("one network cut into %d column strips, halo exchange over NCCL every step" % world) if partitioned else
                                ("%d independent copies" % world if world > 1 else "single GPU"),
                   "l2": "state arrays (%.1f GB) are far larger than the 126 MB L2, no flush needed" % (nV * 1700 / 1e9),
                   "engine_flags": describe_flags(flags), "teleports": int(r2.teleports), "collisions": int(r2.collisions), "spinup_steps": SPINUP_STEPS},
        "clocks": clocks,
        "e2e": {"value": e2e_updates / e2e_s, "unit": "veh*steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": d2h,
                "note": "per step: sumob200_step(1) + sumob200_fcd_begin/_end = FCD rows (48 B: slot, lane, pos, speed, x, y, angle) of every "
                        "on-road vehicle compacted on the device and written to pinned host memory on a copy stream, overlapping the next step; "
                        "the scenario is uploaded once at create, a step has no host input"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": "k_plan", "achieved": plan_gbs, "peak": peak, "unit": "GB/s", "frac": plan_gbs / peak,
                     "traffic": traffic, "traffic_source": traffic_src,
                     "peak_source": peak_src, "bytes_per_update": BYTES_PER_UPDATE_PLAN, "kernel_ms": plan_ms,
                     "window": "%d steps after the timed region; step time there %.4f ms" % (PROF_STEPS, prof_ms / PROF_STEPS),
                     "phase_ms_per_step": {k: v / PROF_STEPS for k, v in phases.items()},
                     "whole_step": {"bytes_per_update": BYTES_PER_UPDATE_STEP, "achieved": BYTES_PER_UPDATE_STEP * value / 1e9,
                                    "frac": BYTES_PER_UPDATE_STEP * value / 1e9 / (peak * world)}},
    }
    if xml is not None:
        line["e2e_xml"] = xml
    if halo is not None:
        line["halo"] = halo
    if flags != 0:
        line["invalid"] = "engine raised flags: %s" % ", ".join(describe_flags(flags))
    if world == 1 and not args.no_cpu_baseline:
        cb = reference_ups(SPINUP_STEPS + 180, args.sigma)
        if cb is not None:
            line["cpu_baseline"] = {"value": cb["ups"], "unit": "veh*steps/s", "cores": cb["procs"], "kind": "reference", "sample": reference_sample_text(cb)}
    print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
