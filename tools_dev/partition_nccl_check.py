#!/usr/bin/env python3
"""torchrun worker: a partitioned run (one rank per GPU, NCCL halo exchange) against a reference trace, bit for bit.
usage: torchrun --nproc-per-node N tools_dev/partition_nccl_check.py X.flat X.trace [steps]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def main():
    import torch
    import torch.distributed as dist
    from sumo_b200.flatfile import load_flat, load_trace
    from sumo_b200.partition import LV_AUTH, PartitionedEngine
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    g = load_flat(sys.argv[1])
    ref = load_trace(sys.argv[2])
    if len(sys.argv) > 3:
        ref = ref[:int(sys.argv[3])]
    numid = g["veh_numid"]
    order = np.argsort(numid, kind="stable")
    eng = PartitionedEngine(g, dist, rank, world, device=local)
    vis = eng.plan["lane_vis"]
    # scenario with E2 detectors: rank 0 also runs ONE engine over the whole network; the merged interval records (a collective
    # read-out every 30 steps) must be that engine's, field by field
    one = None
    det_ok, det_rows = True, 0
    if "e2_lane" in g and rank == 0:
        from sumo_b200.engine import Engine
        one = Engine(g, device=local)
    compared = torch.zeros(2, dtype=torch.int64, device="cuda")
    ok = True
    for (t, ids, lanes, pos, speed) in ref:
        r = eng.step(1)
        st = eng.vehicles()[order]
        on = st["lane"] >= 0
        mine = vis[lanes] == LV_AUTH                # the reference's vehicles that are on this rank's lanes
        if not (np.array_equal(numid[order][on], ids[mine]) and np.array_equal(st["lane"][on], lanes[mine])
                and np.array_equal(st["pos"][on].view(np.int64), pos[mine].view(np.int64))
                and np.array_equal(st["speed"][on].view(np.int64), speed[mine].view(np.int64)) and r.flags == 0):
            print("rank %d: mismatch at t=%d (flags 0x%x)" % (rank, t, r.flags), flush=True)
            ok = False
            break
        compared[0] += int(on.sum())
        compared[1] += len(ids)
        if one is not None:
            one.step(1, result=False)
        if "e2_lane" in g and eng.time_ms % 30000 == 0:
            merged = eng.e2_interval(g["e2_lane"])
            if one is not None:
                want = one.e2_interval()
                det_ok = det_ok and all(np.array_equal(want[k], merged[k]) for k in want.dtype.names)
                det_rows += int((want["n_veh_entered"] > 0).sum())
    eng.close()
    if one is not None:
        one.close()
        print("PARTITION-NCCL-E2 %s (%d busy interval records)" % ("OK" if det_ok and det_rows > 100 else "FAILED", det_rows), flush=True)
        ok = ok and det_ok and det_rows > 100
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    dist.all_reduce(compared[:1])
    if rank == 0:
        total = int(compared[0].item())
        if flag.item() == 1 and total == int(compared[1].item()):
            print("PARTITION-NCCL-OK ranks=%d steps=%d vehicle-steps=%d bit-identical to the reference" % (world, len(ref), total), flush=True)
        else:
            print("PARTITION-NCCL-FAILED", flush=True)
    dist.destroy_process_group()
    return 0 if flag.item() == 1 else 1


if __name__ == "__main__":
    sys.exit(main())
