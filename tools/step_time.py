"""Whole-step timing (GPU box): eager vs CUDA-graph VecEnvRLGames.step at several env counts.
argv[1] = "nosc": Humanoid with enable_self_collisions off (the reference default is on, Humanoid.yaml:80)."""
import os, sys, statistics, io, contextlib
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from omniisaacgymenvs_b200.utils.task_util import make_env
NOSC = len(sys.argv) > 1 and sys.argv[1] == "nosc"
for robot in (("Humanoid",) if NOSC else ("Humanoid", "Ant")):
    for n in [int(x) for x in os.environ.get("STEP_NS", "1024,4096,16384,65536").split(",")]:
        out = [f"{robot} n={n:6d}"]
        for graph in (False, True):
            with contextlib.redirect_stdout(io.StringIO()):
                env, task = make_env(robot, num_envs=n, overrides={"task.sim.Humanoid.enable_self_collisions": False} if NOSC else None)
                env.reset()
            if graph: env.enable_cuda_graph()
            dev = torch.device(task.device)
            acts = [torch.rand(n, task.num_actions, device=dev) * 2 - 1 for _ in range(8)]
            for i in range(20): env.step(acts[i % 8])
            torch.cuda.synchronize()
            import time
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter(); e0.record()
            for i in range(200): env.step(acts[i % 8])
            e1.record(); torch.cuda.synchronize(); wall = (time.perf_counter() - t0) / 200
            ms = e0.elapsed_time(e1) / 200
            out.append(f"{'graph' if graph else 'eager'}: {ms*1e3:7.1f} us/step dev, {wall*1e6:7.1f} us wall ({n/ms/1e3:6.1f} M env-steps/s)")
            del env, task
        print(" | ".join(out), flush=True)
