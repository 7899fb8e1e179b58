"""BASELINE.json configs[4]: Ant + Humanoid (+ Anymal) env-count sweep, 16-step random-policy rollouts.

    python tools/env_sweep.py                                  # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 --master-port 29511 \
        tools/env_sweep.py                                     # G GPUs, env-index sharding

For each task and each TOTAL env count N in 1k..1M the N envs are sharded by index over the ranks; every rank runs
W warm-up rollouts and R timed rollouts of 16 ``VecEnvRLGames.step`` calls (CUDA-graph mode, random-policy actions
drawn on the device, resets happening organically), each rollout bracketed by CUDA events on the launch stream.
The NCCL all-reduce of the episode statistics is issued once per rollout, as in bench.py.  Reported: median rollout
time (max over ranks), whole-job env-steps/s and us per step.  One line per (task, N); rank 0 prints.
"""
import argparse
import contextlib
import io
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from omniisaacgymenvs_b200.utils.distributed import EpisodeStatsReducer, max_over_ranks, shard_range
from omniisaacgymenvs_b200.utils.task_util import make_env

ROLLOUT = 16


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tasks", default="Humanoid,Ant,Anymal")
    ap.add_argument("--counts", default="1024,4096,16384,65536,262144,1048576")
    ap.add_argument("--rollouts", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--eager", action="store_true", help="time the 3-launch eager step instead of the CUDA graph")
    ap.add_argument("--no-self-collision", action="store_true", help="Humanoid with enable_self_collisions off")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    if rank == 0:
        print(f"# {torch.cuda.get_device_name(local)} x{world}; {ROLLOUT}-step random-policy rollouts, "
              f"{'eager' if args.eager else 'CUDA-graph'} step, median of {args.rollouts} rollouts, max over ranks")
        print(f"# {'task':9s} {'total envs':>10s} {'envs/GPU':>9s} {'rollout ms':>11s} {'us/step':>9s} {'M env-steps/s':>14s}")
    for task_name in args.tasks.split(","):
        for total in (int(c) for c in args.counts.split(",")):
            lo, n = shard_range(total, rank, world)
            with contextlib.redirect_stdout(io.StringIO()):
                ov = {"task.sim.Humanoid.enable_self_collisions": False} if (args.no_self_collision and task_name == "Humanoid") else None
                env, task = make_env(task_name, num_envs=n, device_id=local, seed=42, env_id_offset=lo, global_num_envs=total,
                                     overrides=ov)
                env.reset()
                if not args.eager:
                    env.enable_cuda_graph()
            reducer = EpisodeStatsReducer(dev)
            gen = torch.Generator(device=dev).manual_seed(7 + rank)
            A = task.num_actions
            times = []
            for r in range(args.warmup + args.rollouts):
                if dist is not None:
                    dist.barrier()
                torch.cuda.synchronize(dev)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(ROLLOUT):
                    env.step(torch.rand(n, A, device=dev, generator=gen) * 2 - 1)
                e1.record()
                reducer.reduce(task.read_episode_stats())
                torch.cuda.synchronize(dev)
                if r >= args.warmup:
                    times.append(max_over_ranks(e0.elapsed_time(e1), dev))
            times.sort()
            ms = times[len(times) // 2]
            if rank == 0:
                print(f"  {task_name:9s} {total:10d} {n:9d} {ms:11.3f} {ms * 1e3 / ROLLOUT:9.1f} {total * ROLLOUT / ms / 1e3:14.1f}",
                      flush=True)
            del env, task
            torch.cuda.empty_cache()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
