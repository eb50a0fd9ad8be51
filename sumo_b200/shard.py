"""Multi-GPU sharding of a district workload: one process per GPU, every rank owns a contiguous range of independent
districts (sub-networks that share no edge, so the step has no data-path exchange); the only collective is the reduction
of the per-rank counters and the max-over-ranks time ("units are edges; interactions are local").
torch.distributed is plumbing here (NCCL on the GPU box, gloo in the CPU tests); the engine itself never sees it."""


def district_range(total_districts, rank, world):
    """contiguous, balanced split: the first (total % world) ranks get one district more"""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    q, r = divmod(total_districts, world)
    begin = rank * q + min(rank, r)
    return begin, begin + q + (1 if rank < r else 0)


def reduce_step_metrics(dist, device, elapsed_ms, vehicle_updates):
    """whole-job numbers from per-rank ones: (max over ranks of the device time, sum over ranks of the vehicle updates).
    `dist` = torch.distributed (already initialised) or None for a single process."""
    if dist is None:
        return elapsed_ms, vehicle_updates
    import torch
    t = torch.tensor([float(elapsed_ms)], dtype=torch.float64, device=device)
    u = torch.tensor([float(vehicle_updates)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(u, op=dist.ReduceOp.SUM)
    return t.item(), u.item()
