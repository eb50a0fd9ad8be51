"""CPU suite: the N>1 host logic (district sharding, max-over-ranks time, summed updates) with world_size 2 over gloo."""
import os

import pytest

from sumo_b200.shard import district_range, reduce_step_metrics


def test_district_range_is_a_partition():
    for total in (0, 1, 7, 556, 5560):
        for world in (1, 2, 3, 4, 8):
            seen = []
            for rank in range(world):
                b, e = district_range(total, rank, world)
                assert 0 <= b <= e <= total
                seen += list(range(b, e))
            assert seen == list(range(total))
            sizes = [district_range(total, r, world)[1] - district_range(total, r, world)[0] for r in range(world)]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        district_range(10, 2, 2)


def _worker(rank, world, port, out):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    b, e = district_range(11, rank, world)
    ms, updates = reduce_step_metrics(dist, torch.device("cpu"), 10.0 + rank, (e - b) * 1800)
    out[rank] = (ms, updates)
    dist.barrier()
    dist.destroy_process_group()


def test_reduce_step_metrics_gloo_world2():
    import torch.multiprocessing as mp
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(2, 29631, out), nprocs=2, join=True)
    assert out[0] == out[1] == (11.0, 11 * 1800.0)


def test_single_process_passthrough():
    assert reduce_step_metrics(None, None, 3.5, 42) == (3.5, 42)
