"""Cartpole measurements (GPU box): (1) the obs/reward/done pass alone at 1 Mi..16 Mi envs against its
algorithmic bytes (24 B read + 36 B written = 60 B/env-step), (2) the whole VecEnvRLGames.step (reset+efforts ->
K2 x2 substeps -> obs/reward/done), eager and as a CUDA graph, at the reference's 512 envs and larger counts,
(3) the CPU oracle (torch fp32 restatement of tasks/cartpole.py + float64 dynamics oracle) on the host cores.

    python tests/cartpole_time.py > gpurun_out/cartpole_time.txt
"""
import contextlib
import ctypes as C
import io
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

from omniisaacgymenvs_b200 import _lib
from omniisaacgymenvs_b200.utils.task_util import make_env

DEV = torch.device("cuda:0")
PEAKS = os.path.join(ROOT, "MEASURED_PEAKS.json")


def hbm_peak():
    try:
        d = json.load(open(PEAKS))
        for k in ("hbm_gbs", "hbm_gbps"):
            if k in d:
                return float(d[k]), k
    except Exception:
        pass
    return 6551.0, "fallback (B200_PROFILING.md)"


def time_kernel(fn, iters=50, flush=None):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.add_(1.0)  # > L2 write between timed launches
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))


def post_pass():
    lib = _lib.load()
    peak, src = hbm_peak()
    print(f"# obs/reward/done pass, 60 B/env-step algorithmic; HBM peak {peak:.0f} GB/s ({src})")
    flush = torch.zeros(64 << 20, device=DEV)  # 256 MiB > 126 MB L2
    for n in (1 << 20, 1 << 22, 1 << 24):
        q = torch.randn(n, 2, device=DEV); qd = torch.randn(n, 2, device=DEV)
        obs = torch.zeros(n, 4, device=DEV); rew = torch.zeros(n, device=DEV)
        rst = torch.zeros(n, dtype=torch.int64, device=DEV); prog = torch.zeros(n, dtype=torch.int64, device=DEV)
        st = _lib.OigeState(None, None, None, q.data_ptr(), qd.data_ptr(), None, None)
        bf = _lib.OigeCartpoleBuffers(obs.data_ptr(), rew.data_ptr(), rst.data_ptr(), prog.data_ptr(), None, None, None)
        p = _lib.OigeCartpoleParams(0, 1, 1, 0, 3.0, 400.0, 500.0, 5.0)
        s = _lib.current_stream_ptr(DEV)
        ms = time_kernel(lambda: lib.oige_cartpole_obs_reward_reset_f32(C.byref(p), C.byref(st), C.byref(bf), n, s), flush=flush)
        gbs = 60.0 * n / ms / 1e6
        print(f"n={n:9d}  {ms*1e3:8.1f} us  {gbs:7.0f} GB/s  frac {gbs/peak:.2f}  {n/ms/1e6:7.2f} G env-steps/s")


def whole_step():
    print("# whole step (3 launches; K2 = 2 substeps on the 3-link tree)")
    for n in (512, 4096, 16384, 65536, 262144, 1 << 20):
        row = [f"n={n:8d}"]
        for graph in (False, True):
            with contextlib.redirect_stdout(io.StringIO()):
                env, task = make_env("Cartpole", num_envs=n)
                env.reset()
            if graph:
                env.enable_cuda_graph()
            acts = [torch.rand(n, 1, device=DEV) * 2 - 1 for _ in range(8)]
            for i in range(20):
                env.step(acts[i % 8])
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter(); e0.record()
            for i in range(200):
                env.step(acts[i % 8])
            e1.record(); torch.cuda.synchronize()
            wall = (time.perf_counter() - t0) / 200
            ms = e0.elapsed_time(e1) / 200
            row.append(f"{'graph' if graph else 'eager'} {ms*1e3:7.1f} us dev {wall*1e6:7.1f} us wall {n/ms/1e3:8.1f} M env-steps/s")
            del env, task
        print(" | ".join(row), flush=True)


def cpu_oracle():
    from omniisaacgymenvs_b200.robots.model import build_model
    from oracle import cartpole_oracle as co
    from oracle import dynamics as dyn

    n, steps = 16384, 20
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    model = build_model("Cartpole")
    desc = dyn.make_desc(model)
    rng = np.random.default_rng(0)
    pos = np.zeros((n, 3)); pos[:, 2] = 2.0
    quat = np.zeros((n, 4)); quat[:, 0] = 1.0
    vel = np.zeros((n, 6))
    q = rng.uniform(-0.3, 0.3, (n, 2)); qd = rng.normal(size=(n, 2)) * 0.3
    progress = torch.zeros(n, dtype=torch.int64)
    t0 = time.perf_counter()
    for _ in range(steps):
        a = torch.rand(n, 1) * 2 - 1
        eff = co.efforts(co.CARTPOLE, a).double().numpy()
        dyn.step(desc, pos, quat, vel, q, qd, np.ascontiguousarray(eff), None, 2, 0.0083, threads)
        obs, rew, reset, progress = co.post_physics_step(co.CARTPOLE, torch.from_numpy(q).float(), torch.from_numpy(qd).float(), progress)
        ids = reset.nonzero(as_tuple=False).squeeze(-1)
        if len(ids):
            rq, rv = co.reset_values(torch.rand(len(ids), 4))
            q[ids.numpy()] = rq.double().numpy(); qd[ids.numpy()] = rv.double().numpy()
            progress[ids] = 0
    dt = (time.perf_counter() - t0) / steps
    print(f"# CPU oracle port, {threads} threads, n={n}: {dt*1e3:.2f} ms/step = {n/dt/1e6:.2f} M env-steps/s")


if __name__ == "__main__":
    print(torch.cuda.get_device_name(0))
    post_pass()
    whole_step()
    cpu_oracle()
