#!/usr/bin/env python3
"""Throughput benchmark of the step engine: vehicle-updates/s (SUMO's UPS numerator, MSNet.cpp:959) on the synthetic grid.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--vehicles V] [--sigma S]      our arm (libsumob200.so, C ABI)
  python bench.py --impl reference ...                                               the reference's own CPU path

Workload (BASELINE.json configs[1]: grid with static traffic lights, 1M vehicles, Krauss sigma=0.5, single B200): V resident
vehicles on ceil(V/1800) disjoint districts, each the 10x10 / 200 m / traffic_light grid netgenerate builds with 1800
random-walk vehicles (25 veh/lane-km), tiled in array space from tests/golden/bench_grid10_tls.flat (sumo_b200/tiler.py).
A "step" is one simulation step (DELTA_T = 1 s) of all districts.  Under torchrun every rank simulates its own V vehicles
(weak scaling, districts are independent: no data-path collective).
"""
import argparse
import json
import os
import re
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BASE = os.path.join(ROOT, "tests", "golden", "bench_grid10_tls")
VEH_PER_TILE = 1800
TLS_CYCLE_STEPS = 90        # every traffic light of the workload runs the same 90 s program with offset 0: the cost of a step is periodic in it
PROF_STEPS = TLS_CYCLE_STEPS  # window of the per-kernel (roofline) timing: one whole cycle
SPINUP_STEPS = 420          # all departures happen in [0, 300 + 16) s; afterwards the resident count is constant
# algorithmic bytes per vehicle-update (SURVEY.md section 8d / DESIGN.md "roofline"): whole step, and the plan kernel alone
BYTES_PER_UPDATE_STEP = 576
BYTES_PER_UPDATE_PLAN = 320
NCU_PLAN_TRAFFIC_BYTES = 3733032000 + 1453889000    # profiles/r01_k_plan_ncu.txt: dram__bytes_read.sum + dram__bytes_write.sum of one k_plan launch (mostly register-spill traffic)


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region (B200_PROFILING.md's clocks line).  NVML is polled from a
    thread every 5 ms (the timed region of the default run is ~0.15 s, shorter than nvidia-smi's start-up); nvidia-smi -lms is
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


def reference_ups(total_steps, procs=None, repeats=3):
    """the reference's own CPU implementation (oracle/_ref/sumo, unmodified sources).  `sumo --threads` needs the FOX toolkit
    (src/microsim/MSFrame.cpp:1119-1124), which is absent, so one sumo process is single-threaded; the workload is made of
    independent districts, so all host cores are used the way the workload allows: `procs` concurrent sumo processes, one
    district each (the same district file), `repeats` rounds.  Aggregate = sum of the processes' own UPS (vehicle updates /
    their simulation wall time, MSNet.cpp:959) while they run side by side."""
    exe = os.path.join(ROOT, "oracle", "_ref", "sumo")
    if not os.path.exists(exe):
        return None
    procs = procs or max(1, os.cpu_count() or 1)
    cmd = [exe, "-n", BASE + ".net.xml", "-r", BASE + ".rou.xml", "--no-step-log", "--no-warnings", "--time-to-teleport", "-1",
           "--default.departspeed", "0", "--default.departlane", "first", "--end", str(total_steps), "--duration-log.statistics"]
    ups_rounds, dur = [], []
    t0 = time.time()
    for _ in range(repeats):
        ps = [subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for _ in range(procs)]
        total = 0.0
        for p in ps:
            out = p.communicate()[0]
            m = re.search(r"UPS: ([0-9.]+)", out)
            d = re.search(r"Duration: ([0-9.]+)s", out)
            if not m:
                return None
            total += float(m.group(1))
            dur.append(float(d.group(1)) if d else 0.0)
        ups_rounds.append(total)
    wall = time.time() - t0
    return {"ups": float(np.mean(ups_rounds)), "ups_per_process": float(np.mean(ups_rounds)) / procs, "duration_s": float(np.mean(dur)),
            "wall_s": wall, "steps": total_steps, "procs": procs, "repeats": repeats}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    total = SPINUP_STEPS + args.warmup + max(args.steps, 200)
    r = reference_ups(total)
    if r is None:
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/sumo not built"}))
        return 0
    sample = ("%d concurrent sumo processes (one per host core; --threads needs FOX, absent), each 1 district (10x10 TLS grid, %d vehicles) x %d steps, "
              "%d rounds, %.1f s wall; %.0f veh*steps/s per process" % (r["procs"], VEH_PER_TILE, total, r["repeats"], r["wall_s"], r["ups_per_process"]))
    line = {
        "impl": "reference", "metric": "vehicle-updates/s", "value": r["ups"], "unit": "veh*steps/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 * r["duration_s"] / total, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "netgenerate 10x10 grid, 200 m, static traffic lights, Krauss sigma=%.1f, step-length 1 s; reference sumo, one district per process" % args.sigma,
                   "vehicles": VEH_PER_TILE, "processes": r["procs"]},
        "cpu_baseline": {"value": r["ups"], "unit": "veh*steps/s", "cores": r["procs"], "kind": "reference", "sample": sample},
        "e2e": {"value": r["ups"], "unit": "veh*steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=TLS_CYCLE_STEPS)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--vehicles", type=int, default=1000000, help="resident vehicles per GPU")
    ap.add_argument("--sigma", type=float, default=0.5)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--emissions", action="store_true", help="every vehicle carries the HBEFA emission device (BASELINE.json configs[4] flavour)")
    args = ap.parse_args()
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
    from sumo_b200.flatfile import load_flat
    from sumo_b200.tiler import tile

    base = load_flat(BASE + ".flat")
    tiles = max(1, (args.vehicles + VEH_PER_TILE - 1) // VEH_PER_TILE)
    scenario = tile(base, tiles, overrides={"vt_sigma": np.full(base["vt_sigma"].shape, args.sigma)})
    if not args.emissions:
        scenario["veh_emissions"] = np.zeros_like(scenario["veh_emissions"])
    eng = Engine(scenario, device=local_rank)
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
        checksum += float(rows["speed"][-1])
        eng.fcd_begin()
    rows = eng.fcd_end()
    d2h += rows.nbytes + 4
    rr = eng.step(0)
    e2e_s = time.perf_counter() - t0
    e2e_updates = rr.vehicle_updates - r1.vehicle_updates
    d2h = d2h // e2e_steps
    state = eng.vehicles().copy()
    # ---- roofline of the dominant kernel (k_plan): per-phase CUDA events on the engine's stream
    eng.set_profiling(True)
    eng.phase_times(reset=True)
    r2 = eng.step(PROF_STEPS)
    phases = eng.phase_times(reset=True)
    eng.set_profiling(False)
    prof_updates = r2.vehicle_updates - rr.vehicle_updates
    flags = r2.flags
    mean_speed = float(state["speed"][state["lane"] >= 0].mean())

    if dist is not None:
        import torch
        from sumo_b200.shard import reduce_step_metrics
        dev = torch.device("cuda", local_rank)
        ms, updates = reduce_step_metrics(dist, dev, ms, updates)              # max over ranks of the device time, summed updates
        e2e_s, e2e_updates = reduce_step_metrics(dist, dev, e2e_s, e2e_updates)
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return 0

    peak, peak_src = measured_peak_gbs()
    value = updates / (ms / 1000.0)
    plan_ms = phases["plan"] / PROF_STEPS
    plan_gbs = BYTES_PER_UPDATE_PLAN * (prof_updates / PROF_STEPS) / (plan_ms / 1000.0) / 1e9
    line = {
        "metric": "vehicle-updates/s", "value": value, "unit": "veh*steps/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "%d disjoint districts per GPU, each netgenerate 10x10 grid / 200 m / static traffic lights with %d resident vehicles "
                               "(25 veh/lane-km), Krauss sigma=%.1f, step-length 1 s%s" % (tiles, VEH_PER_TILE, args.sigma, ", HBEFA4 emission device on every vehicle" if args.emissions else ""),
                   "vehicles_per_gpu": nV, "resident_per_gpu": int(r1.running), "districts_per_gpu": tiles, "mean_speed_mps": mean_speed,
                   "l2": "state arrays (%.1f GB) are far larger than the 126 MB L2, no flush needed" % (nV * 1700 / 1e9),
                   "engine_flags": describe_flags(flags), "spinup_steps": SPINUP_STEPS},
        "clocks": clocks,
        "e2e": {"value": e2e_updates / e2e_s, "unit": "veh*steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": d2h,
                "note": "per step: sumob200_step(1) + sumob200_fcd_begin/_end = FCD rows (48 B: slot, lane, pos, speed, x, y, angle) of every "
                        "on-road vehicle compacted on the device and copied to pinned host memory on a copy stream, overlapping the next step; "
                        "the scenario is uploaded once at create, a step has no host input"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": "k_plan", "achieved": plan_gbs, "peak": peak, "unit": "GB/s", "frac": plan_gbs / peak,
                     "traffic": NCU_PLAN_TRAFFIC_BYTES if nV == 1000800 else None, "traffic_source": "profiles/r01_k_plan_ncu.txt (dram__bytes_read.sum + dram__bytes_write.sum, one launch, 1,000,800 vehicles)",
                     "peak_source": peak_src, "bytes_per_update": BYTES_PER_UPDATE_PLAN, "kernel_ms": plan_ms,
                     "phase_ms_per_step": {k: v / PROF_STEPS for k, v in phases.items()},
                     "whole_step": {"bytes_per_update": BYTES_PER_UPDATE_STEP, "achieved": BYTES_PER_UPDATE_STEP * value / 1e9,
                                    "frac": BYTES_PER_UPDATE_STEP * value / 1e9 / (peak * world)}},
    }
    if world == 1 and not args.no_cpu_baseline:
        total = SPINUP_STEPS + 280
        cb = reference_ups(total)
        if cb is not None:
            line["cpu_baseline"] = {"value": cb["ups"], "unit": "veh*steps/s", "cores": cb["procs"], "kind": "reference",
                                    "sample": "oracle/_ref/sumo (unmodified reference; single-threaded: --threads needs the FOX toolkit, absent): %d concurrent "
                                              "processes, one per host core, each 1 district (%d vehicles) x %d steps, %d rounds, %.1f s wall; %.0f veh*steps/s per process"
                                              % (cb["procs"], VEH_PER_TILE, total, cb["repeats"], cb["wall_s"], cb["ups_per_process"])}
    print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
