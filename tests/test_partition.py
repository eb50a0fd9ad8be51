"""Edge-partitioned multi-GPU run (sumo_b200/partition.py + csrc/sb2_halo.h).

CPU suite: the strip planning, the transport with world_size 2 over gloo, and -- through tests/host_emul, which steps all strips in
one process with the very pack / unpack functions the device kernels are made of -- that a run cut into 2 or 3 strips reproduces
the REFERENCE's trace bit for bit (so it equals the single-strip run as well), including teleports that cross the cut.
GPU suite: the same on the device, all strips as engines of one process on one GPU (LocalPartitionedRun), and over NCCL with one
process per GPU when the box has two."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT, reference_case
from sumo_b200 import partition
from sumo_b200.flatfile import load_flat, load_trace
from test_parity_host import assert_traces_identical


def test_plan_is_a_partition_and_lists_mirror():
    flat, _ = reference_case("c1_grid10")
    g = load_flat(flat)
    for world in (1, 2, 3, 5):
        plans = [partition.plan_partition(g, r, world) for r in range(world)]
        own = np.stack([p["lane_vis"] == partition.LV_AUTH for p in plans])
        assert (own.sum(axis=0) == 1).all()              # every lane has exactly one owner
        for r, p in enumerate(plans):
            vis = p["lane_vis"]
            assert (vis[p["exp"][0]] == partition.LV_AUTH).all() and (vis[p["exp"][1]] == partition.LV_AUTH).all()
            assert set(vis[np.concatenate(p["imp"])]) <= {partition.LV_GHOST, partition.LV_HALO}
            # whatever is visible but not owned is imported from exactly one side
            seen = np.zeros(len(vis), int)
            for side in (0, 1):
                seen[p["imp"][side]] += 1
            assert ((seen == 1) == ((vis == partition.LV_GHOST) | (vis == partition.LV_HALO))).all()
            if r + 1 < world:
                assert np.array_equal(p["exp"][1], plans[r + 1]["imp"][0]) and np.array_equal(p["imp"][1], plans[r + 1]["exp"][0])
            assert p["capacity"] == plans[0]["capacity"]
    with pytest.raises(ValueError):
        partition.plan_partition(g, 0, 6)               # 10 columns: at most 5 strips of 2


@pytest.mark.parametrize("case,world", [("c1_grid10", 2), ("c1_grid10", 3), ("grid5_dense", 2)])
def test_partitioned_emulation_matches_reference(case, world, host_emul, tmp_path):
    """grid5_dense: 51 teleports until t = 900 s (reference default time-to-teleport), several of them across the cut"""
    flat, trace = reference_case(case)
    part = str(tmp_path / "p.part")
    partition.write_partition_file(part, load_flat(flat), world)
    out = str(tmp_path / "p.trace")
    ref = load_trace(trace)
    r = subprocess.run([host_emul, flat, "--partitions", part, "--trace-out", out, "--steps", str(len(ref))], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "errors=0" in r.stderr
    assert assert_traces_identical(ref, load_trace(out)) > 1000


@pytest.mark.parametrize("case,world", [("c1_grid10", 2), ("c1_grid10", 3), ("grid5_dense", 2)])
def test_partitioned_emulation_reports_the_detectors_of_the_single_run(case, world, host_emul, tmp_path):
    """E1 and E2 detectors under the edge partition: a detector is reported by the partition that owns its lane (vehicles of the
    ghost column are executed redundantly; reminders are forgotten when a vehicle is only read, rebuilt for a teleporting one).
    The merged interval lines must be the single run's -- which tests/test_detectors.py / test_e2_detectors.py pin to the
    reference's files -- character by character."""
    from test_detectors import detector_case
    from test_e2_detectors import e2_case
    for kind, (flat, trace, _), flag in (("e1", detector_case(case), "--e1"), ("e2", e2_case(case), "--e2")):
        n = len(load_trace(trace))
        part = str(tmp_path / (kind + ".part"))
        partition.write_partition_file(part, load_flat(flat), world)
        one, many = str(tmp_path / (kind + ".one.txt")), str(tmp_path / (kind + ".many.txt"))
        subprocess.run([host_emul, flat, flag + "-out", one, flag + "-period", "30000", "--steps", str(n)], check=True, timeout=900, stderr=subprocess.DEVNULL)
        r = subprocess.run([host_emul, flat, "--partitions", part, flag + "-out", many, flag + "-period", "30000", "--steps", str(n)],
                           capture_output=True, text=True, timeout=900)
        assert r.returncode == 0 and "errors=0" in r.stderr, r.stderr[-2000:]
        a = open(one).read().splitlines()
        b = open(many).read().splitlines()
        # (the single run also writes the partial interval at the end of the simulation)
        assert len(b) > 200 and a[:len(b)] == b, next((x, y) for x, y in zip(a, b) if x != y)


def _worker(rank, world, port, out):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    send = [torch.full((64,), 10 * rank + side, dtype=torch.uint8) for side in (0, 1)]
    recv = [torch.zeros(64, dtype=torch.uint8) for _ in (0, 1)]
    for _ in range(3):
        partition.exchange_messages(dist, rank, world, send, recv)
    out[rank] = (int(recv[0][0]), int(recv[1][0]))
    dist.barrier()
    dist.destroy_process_group()


def test_exchange_messages_gloo_world2():
    import torch.multiprocessing as mp
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(2, 29641, out), nprocs=2, join=True)
    # rank 0 has only a right neighbour (its side 1 receives rank 1's side 0 message), rank 1 only a left one
    assert out[0] == (0, 10) and out[1] == (1, 0)


@pytest.mark.gpu
@pytest.mark.parametrize("case,world", [("c1_grid10", 2), ("c1_grid10", 3), ("grid5_dense", 2)])
def test_partitioned_engines_match_reference_on_one_gpu(case, world):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs CUDA")
    flat, trace = reference_case(case)
    g = load_flat(flat)
    ref = load_trace(trace)
    numid = g["veh_numid"]
    order = np.argsort(numid, kind="stable")
    run = partition.LocalPartitionedRun(g, world)
    compared = 0
    try:
        for (t, ids, lanes, pos, speed) in ref:
            run.step()
            st = run.vehicles()[order]
            on = st["lane"] >= 0
            assert np.array_equal(numid[order][on], ids), "t=%d: sets of vehicles on the road differ" % t
            assert np.array_equal(st["lane"][on], lanes), "t=%d: lanes differ" % t
            assert np.array_equal(st["pos"][on].view(np.int64), pos.view(np.int64)), "t=%d: positions differ (bitwise)" % t
            assert np.array_equal(st["speed"][on].view(np.int64), speed.view(np.int64)), "t=%d: speeds differ (bitwise)" % t
            compared += len(ids)
        res = run.results()
        assert all(r.flags == 0 for r in res)
        assert sum(r.running for r in res) == len(ref[-1][1]) + sum(r.teleporting for r in res)
    finally:
        run.close()
    assert compared > 1000


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 3])
def test_partitioned_edgedata_sums_to_the_single_engine(world):
    """edgeData accumulators (sb2_meandata.h) under the edge partition: every rank reports its OWN lanes only (the vehicles of its
    ghost column are executed redundantly, their lanes are the neighbour's); merged by lane they equal one engine over the whole
    network -- counters exactly, the real sums to 1e-9 (atomics: order of additions)."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs CUDA")
    from test_meandata import meandata_case
    from sumo_b200.engine import Engine
    flat, trace, _ = meandata_case("c1_grid10")
    g = load_flat(flat)
    steps = 900
    with Engine(flat) as one:
        one.step(steps, result=False)
        want = one.meandata_interval().copy()
    run = partition.LocalPartitionedRun(g, world)
    try:
        for _ in range(steps):
            run.step()
        merged = np.zeros_like(want)
        seen = np.zeros(len(want), int)
        for p in run.parts:
            md = p["eng"].meandata_interval()
            own = p["plan"]["lane_vis"] == partition.LV_AUTH
            assert (md["sample_seconds"][~own] == 0).all() and (md["entered"][~own] == 0).all() and (md["departed"][~own] == 0).all()
            merged[own] = md[own]
            seen[own] += 1
        assert (seen == 1).all()
    finally:
        run.close()
    for k in ("departed", "arrived", "entered", "left", "teleported"):
        assert np.array_equal(merged[k], want[k]), k
    assert want["entered"].sum() > 2000 and want["arrived"].sum() > 100
    for k in ("sample_seconds", "travelled_distance", "wait_seconds", "time_loss", "front_sample_seconds", "front_travelled_distance",
              "veh_length_sum", "occupation_sum", "minimal_vehicle_length"):
        assert np.allclose(merged[k], want[k], rtol=1e-9, atol=1e-9), k


@pytest.mark.gpu
@pytest.mark.parametrize("case,world", [("c1_grid10", 3), ("grid5_dense", 2)])
def test_partitioned_detectors_equal_the_single_engine(case, world):
    """E1 and E2 detectors under the edge partition on the device: every rank reports the detectors on its OWN lanes; merged by
    owner the interval records equal one engine over the whole network (the E2 sums bit for bit: they are added per detector in
    sorted order; E1 means are reduced from the same events)."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs CUDA")
    from test_detectors import detector_case
    from test_e2_detectors import e2_case
    from sumo_b200.engine import Engine
    for kind, (flat, trace, _) in (("e1", detector_case(case)), ("e2", e2_case(case))):
        g = load_flat(flat)
        steps = (len(load_trace(trace)) // 30) * 30
        det_lane = g[kind + "_lane"]

        def read(eng):
            return eng.e1_interval(eng.time_ms - 30000, eng.time_ms).copy() if kind == "e1" else eng.e2_interval().copy()
        want = []
        with Engine(flat) as one:
            for _ in range(steps // 30):
                one.step(30, result=False)
                want.append(read(one))
        run = partition.LocalPartitionedRun(g, world)
        try:
            got = []
            for k in range(steps):
                run.step()
                if (k + 1) % 30 == 0:
                    merged = np.zeros_like(want[0])
                    seen = np.zeros(len(det_lane), int)
                    for p in run.parts:
                        rows = read(p["eng"])
                        own = p["plan"]["lane_vis"][det_lane] == partition.LV_AUTH
                        merged[own] = rows[own]
                        seen[own] += 1
                    assert (seen == 1).all()
                    got.append(merged)
        finally:
            run.close()
        busy = 0
        for a, b in zip(want, got):
            for name in a.dtype.names:
                assert np.array_equal(a[name], b[name]), (kind, name)
            busy += int((a["n_veh_entered"] > 0).sum())
        assert busy > 200


@pytest.mark.gpu
def test_two_gpus_over_nccl_match_reference():
    """one process per GPU (torchrun), halo exchange over NCCL: tools_dev/partition_nccl_check.py compares every rank's own
    vehicles with the reference trace after every step"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from test_e2_detectors import e2_case
    flat, trace, _ = e2_case("c1_grid10")      # (with E2 detectors: the tool also checks the collective interval read-out)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29655", os.path.join(ROOT, "tools_dev", "partition_nccl_check.py"), flat, trace],
                       capture_output=True, text=True, timeout=1200)
    assert r.returncode == 0, (r.stdout + r.stderr)[-3000:]
    assert "PARTITION-NCCL-OK" in r.stdout and "PARTITION-NCCL-E2 OK" in r.stdout
