"""Env-index sharding across the GPUs of one box.

Environments never interact (the reference filters cross-env collisions, tasks/base/rl_task.py:
127-128), so the step needs NO per-step collective: rank r owns the contiguous global env range
``shard_range(N, r, G)``, env origins and reset noise are keyed by the GLOBAL env id so results do
not depend on G.  The only exchange is an all-reduce (NCCL over NVLink on GPUs, gloo in the CPU
tests) of the tiny episode-statistics vector every K steps.
"""
from typing import Optional, Tuple

import torch


def shard_range(global_num_envs: int, rank: int, world_size: int) -> Tuple[int, int]:
    """(offset, count) of the envs rank owns; shards are contiguous, sizes differ by at most one tile
    of 32 envs so that every shard starts on a K1 tile boundary."""
    if world_size < 1 or not 0 <= rank < world_size:
        raise ValueError(f"bad rank/world_size {rank}/{world_size}")
    tiles = (global_num_envs + 31) // 32
    base, extra = divmod(tiles, world_size)
    start_tile = rank * base + min(rank, extra)
    n_tiles = base + (1 if rank < extra else 0)
    start = min(start_tile * 32, global_num_envs)
    end = min((start_tile + n_tiles) * 32, global_num_envs)
    return start, end - start


class EpisodeStatsReducer:
    """Accumulates [sum reward, episodes finished, sum episode length, env-steps] over ranks.

    ``reduce(local)`` all-reduces ``local`` (a [4] float64 tensor on the step's device) on a side
    stream when CUDA is used, so the collective never sits on the env-step stream."""

    def __init__(self, device, group=None):
        self.device = torch.device(device)
        self.group = group
        self.total = torch.zeros(4, dtype=torch.float64, device=self.device)
        self._side: Optional[torch.cuda.Stream] = None
        if self.device.type == "cuda":
            self._side = torch.cuda.Stream(device=self.device)

    @staticmethod
    def _dist():
        import torch.distributed as dist

        return dist if dist.is_available() and dist.is_initialized() else None

    def reduce(self, local: torch.Tensor) -> None:
        dist = self._dist()
        if self._side is not None:
            self._side.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(self._side):
                if dist is not None:
                    dist.all_reduce(local, group=self.group)
                self.total.add_(local)
            local.record_stream(self._side)
        else:
            if dist is not None:
                dist.all_reduce(local, group=self.group)
            self.total.add_(local)

    def result(self) -> dict:
        if self._side is not None:
            torch.cuda.current_stream(self.device).wait_stream(self._side)
        s = self.total.tolist()
        # s[0]: sum of the returns of FINISHED episodes when the task passes OigeTaskBuffers.episode_return (the
        # locomotion tasks do), else the sum of every env-step's reward (Cartpole / Anymal passes)
        return {"sum_reward": s[0], "episodes_finished": s[1],
                "mean_episode_return": (s[0] / s[1]) if s[1] else None,
                "mean_episode_length": (s[2] / s[1]) if s[1] else None, "env_steps": s[3]}


def max_over_ranks(value: float, device) -> float:
    """Device-timed durations are reported as the max over ranks."""
    import torch.distributed as dist

    t = torch.tensor([value], dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized():
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
