"""Post-physics pass (obs / reward / done) of every task at bandwidth sizes, against its algorithmic bytes (GPU box).

    python tools/post_pass_time.py > gpurun_out/post_pass_time.txt

Times ``task.post_physics_step()`` (one launch) with CUDA events on the launch stream, L2 flushed between launches
(256 MiB write), median of 30.  Algorithmic bytes per env-step: Humanoid 772, Ant 556, Anymal 624
(csrc/anymal.cu header), Cartpole 60 (csrc/cartpole.cu header).
"""
import contextlib
import io
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

from omniisaacgymenvs_b200.utils.task_util import make_env

BYTES = {"Humanoid": 772, "Ant": 556, "Anymal": 624, "Cartpole": 60}


def hbm_peak():
    try:
        d = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        for k in ("hbm_gbs", "hbm_gbps"):
            if k in d:
                return float(d[k]), "MEASURED_PEAKS.json " + k
    except Exception:
        pass
    return 6551.0, "fallback"


def main():
    dev = torch.device("cuda:0")
    peak, src = hbm_peak()
    print(f"# {torch.cuda.get_device_name(0)}; HBM peak {peak:.0f} GB/s ({src}); post_physics_step, one launch, L2 flushed")
    flush = torch.zeros(64 << 20, device=dev)
    for name in ("Humanoid", "Ant", "Anymal", "Cartpole"):
        for n in (4096, 16384, 262144, 1 << 20, 1 << 22):
            if name == "Cartpole" and n < (1 << 20):
                continue
            with contextlib.redirect_stdout(io.StringIO()):
                env, task = make_env(name, num_envs=n)
                env.reset()
            for _ in range(4):
                env.step(torch.rand(n, task.num_actions, device=dev) * 2 - 1)
            ts = []
            for i in range(35):
                flush.add_(1.0)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); task.post_physics_step(); e1.record()
                torch.cuda.synchronize()
                if i >= 5:
                    ts.append(e0.elapsed_time(e1))
            ms = float(np.median(ts))
            gbs = BYTES[name] * n / ms / 1e6
            print(f"{name:9s} n={n:8d}  {ms*1e3:8.1f} us  {gbs:7.0f} GB/s  frac {gbs/peak:.2f}  {n/ms/1e6:7.2f} G env/s", flush=True)
            del env, task
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
