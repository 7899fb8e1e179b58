"""CPU, world_size 2 over gloo: the host-side logic of the multi-GPU path (shard ranges, env-origin
slices, episode-statistics all-reduce, max-over-ranks timing).  The GPU side of sharding (bit-exact
shard == rows of the unsharded run) is tests/test_env_gpu.py::test_sharded_envs_equal_rows..."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from omniisaacgymenvs_b200.articulations import grid_env_origins
from omniisaacgymenvs_b200.utils.distributed import EpisodeStatsReducer, max_over_ranks, shard_range


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_global, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        off, cnt = shard_range(n_global, rank, world)
        origins = grid_env_origins(n_global, 5.0, off, cnt)
        red = EpisodeStatsReducer("cpu")
        for step in range(3):  # three reductions of rank-dependent local stats
            local = torch.tensor([1.5 * (rank + 1), float(rank), 10.0 * rank, float(cnt)], dtype=torch.float64)
            red.reduce(local)
        slowest = max_over_ranks(1.0 + rank, "cpu")
        gathered = [None] * world
        dist.all_gather_object(gathered, (off, cnt, float(origins.sum())))
        np.save(os.path.join(out_dir, f"r{rank}.npy"),
                np.array([off, cnt, slowest] + red.total.tolist() + [g[1] for g in gathered], dtype=np.float64))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_global", [65536, 1000])
def test_two_rank_sharding_and_stats_allreduce(tmp_path, n_global):
    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, n_global, str(tmp_path)), nprocs=world, join=True)
    r = [np.load(tmp_path / f"r{k}.npy") for k in range(world)]
    offs, cnts = [int(x[0]) for x in r], [int(x[1]) for x in r]
    assert offs[0] == 0 and offs[1] == cnts[0] and sum(cnts) == n_global
    assert all(o % 32 == 0 for o in offs)
    for x in r:
        assert x[2] == 2.0  # max over ranks of (1 + rank)
        # three all-reduced steps: sums over ranks of the local vectors
        assert np.allclose(x[3:7], [3 * (1.5 + 3.0), 3 * 1.0, 3 * 10.0, 3 * n_global])
        assert [int(v) for v in x[7:9]] == cnts


def test_shard_range_properties():
    for n in (1, 31, 32, 33, 4096, 16384, 65536, 1 << 20, 1000003):
        for g in (1, 2, 3, 4, 8):
            parts = [shard_range(n, r, g) for r in range(g)]
            assert parts[0][0] == 0 and sum(c for _, c in parts) == n
            for (o1, c1), (o2, _) in zip(parts, parts[1:]):
                assert o1 + c1 == o2
            assert all(o % 32 == 0 for o, c in parts if c)
            sizes = [c for _, c in parts if c]
            assert max(sizes) - min(sizes) <= 32 + 31
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def test_grid_slices_concatenate_to_the_full_grid():
    n, g = 65536, 8
    full = grid_env_origins(n, 5.0)
    parts = [grid_env_origins(n, 5.0, *shard_range(n, r, g)) for r in range(g)]
    assert np.array_equal(np.concatenate(parts), full)
