#!/usr/bin/env python
"""Benchmark of the locomotion environment step (BASELINE.json metric: Humanoid env-steps/s at
16k envs/GPU; obs+reward HBM GB/s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--envs-per-gpu E]

A "step" is one VecEnvRLGames.step of every environment: clamp actions, device-side reset of
flagged envs, controlFrequencyInv=2 articulation substeps (self-collision on, as in the reference's
Humanoid.yaml), observations / reward / termination -- ONE launch of the whole-step kernel K2-F
(`env_step_role_kernel`) per step.  One process per GPU; environments are partitioned by index, no
per-step collective; NCCL all-reduces the episode statistics every 16 steps on a side stream.

Printed JSON (rank 0, one line): the driver contract keys plus
  roofline      the dominant kernel (K2-F, the whole step) against the HBM roofline, measured live; it is issue /
                latency-bound, so `issue` (warp-instructions per launch from the committed ncu capture / live time
                against the SMs' issue peak) is the figure that explains it
  roofline_k1   the stand-alone obs/reward pass K1 at 1,048,576 envs (772 MB per launch > 126 MB L2) and at the
                benchmark's own env count
  sustained     the same CUDA-graph step looped for >= 2 s with its own clock record (a loaded-GPU twin of `value`)
  no_self_collision  `value`'s twin with enable_self_collisions off (the round-1 mechanical system)
  cpu_baseline  the oracle (torch CPU obs/reward + float64 OpenMP dynamics) on a bounded sample, rank 0, N=1
  e2e           the same step through `step_host` (pinned HOST buffers, H2D + D2H inside the timed region)
  step_breakdown per-kernel device time of the staged (hook-by-hook) K3 / K2 / K1 path, for reference
`--impl reference` times the CPU oracle port alone (PhysX / Isaac Sim are not installable: see DESIGN.md).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ENVS_PER_GPU = 16384
ROBOT = "Humanoid"
K1_BYTES = {"Humanoid": 772, "Ant": 556}      # SURVEY.md section 8(d) algorithmic bytes per env-step (K1)
STEP_BYTES = {"Humanoid": 944}                # informational: fused full step
ROOFLINE_ENVS = 1 << 20
# dram__bytes_read.sum + dram__bytes_write.sum of one K1 launch at 1,048,576 envs, from the ncu --set full
# capture summarised in profiles/r1_ncu_full.txt (390.1 MB + 367.4 MB); algorithmic bytes are 809.5 MB
K1_TRAFFIC_BYTES_1M = 757.6e6
K1_TRAFFIC_SOURCE = "profiles/r1_ncu_full.txt"
# The dominant kernel, K2-F at 16384 Humanoid envs with self-collision: figures of the committed ncu --set full capture
# (one launch = one whole env step).  Read from profiles/ at run time when the file travels with the repo.
K2F_NCU_FILES = ("profiles/r3_ncu_k2f_sc_16384.txt", "profiles/r2_ncu_k2f_sc_16384.txt")  # newest first
K2F_NCU_DEFAULTS = {"warp_inst": 40.0e6, "issue_util_pct": 42.7, "dram_bytes": 10.3e6, "time_us": 102.6}


def read_k2f_ncu():
    """smsp__inst_executed.sum, issue utilisation, DRAM bytes and duration of the committed K2-F capture."""
    out = dict(K2F_NCU_DEFAULTS)
    out["source"] = "bench.py constants"
    path = next((os.path.join(ROOT, f) for f in K2F_NCU_FILES if os.path.exists(os.path.join(ROOT, f))), "")
    keys = {"smsp__inst_executed.sum": "warp_inst", "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_util_pct",
            "gpu__time_duration.sum": "time_us"}
    try:
        rd = wr = None
        with open(path) as f:
            for ln in f:
                p = ln.split()
                if len(p) >= 3 and p[0] == "#" and p[1] == "source_hash":
                    out["capture_hash"] = p[2]
                if len(p) < 2:
                    continue
                scale = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0}.get(p[2] if len(p) > 2 else "", 1.0)
                if p[0] in keys:
                    out[keys[p[0]]] = float(p[1].replace(",", ""))
                elif p[0] == "dram__bytes_read.sum":
                    rd = float(p[1].replace(",", "")) * scale
                elif p[0] == "dram__bytes_write.sum":
                    wr = float(p[1].replace(",", "")) * scale
        if rd is not None and wr is not None:
            out["dram_bytes"] = rd + wr
        out["source"] = os.path.relpath(path, ROOT)
        # a capture is only as good as the binary it was taken from: tools/recapture_k2f.py heads the file with a hash
        # of the kernel sources; anything else (the round-2 file, or sources edited since) is labelled stale
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        from recapture_k2f import kernel_source_hash
        if out.get("capture_hash") != kernel_source_hash():
            out["source"] += " (stale: kernel sources changed since the capture)"
    except Exception:
        pass
    return out


def read_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            with open(path) as f:
                return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """Samples SM clock and throttle reasons while the benchmark runs (NVML; nvidia-smi as a fallback)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.samples, self._stop, self._t = index, [], threading.Event(), None
        self.sm_max = None
        self.interval = 0.001  # s between NVML samples (raised around the host-buffer loop: one polling thread per rank)
        self._nvml = None
        try:
            import pynvml

            pynvml.nvmlInit()
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(visible.split(",")[index]) if visible and visible.split(",")[index].isdigit() else index
            self._h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.sm_max = float(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
            self._nvml = pynvml
        except Exception:
            self._nvml = None

    def _sample_nvml(self):
        n = self._nvml
        sm = float(n.nvmlDeviceGetClockInfo(self._h, n.NVML_CLOCK_SM))
        try:
            r = int(n.nvmlDeviceGetCurrentClocksEventReasons(self._h))
        except Exception:
            r = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self._h))
        flags = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}
        self.samples.append((sm, [k for k, bit in flags.items() if r & bit]))

    def _sample_smi(self):
        out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i",
                              str(self.index)], capture_output=True, text=True, timeout=5).stdout.strip()
        f = [x.strip() for x in out.split(",")]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        self.sm_max = float(f[1])
        self.samples.append((float(f[0]), [n for i, n in enumerate(names) if f[2 + i].lower().startswith("active")]))

    def _run(self):
        while not self._stop.is_set():
            try:
                self._sample_nvml() if self._nvml is not None else self._sample_smi()
            except Exception:
                pass
            self._stop.wait(self.interval if self._nvml is not None else max(self.interval, 0.1))

    def __enter__(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=6)

    def mark(self):
        return len(self.samples)

    def summary(self, lo=0, hi=None):
        part = self.samples[lo:hi]
        sm = [s[0] for s in part]
        reasons = sorted({r for s in part for r in s[1]})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": self.sm_max, "reasons": reasons,
                "samples": len(part), "source": "nvml" if self._nvml is not None else "nvidia-smi"}


# ------------------------------------------------------------------------------------------------
# CPU reference arm / cpu_baseline: the oracle port on the host cores
# ------------------------------------------------------------------------------------------------
def cpu_env_step_rate(n_envs: int, steps: int, warmup: int, threads: int):
    """Full env step on the CPU: float64 OpenMP dynamics (2 substeps) + torch obs/reward/termination +
    reset of flagged envs.  Returns (env-steps/s, ms per step)."""
    import numpy as np
    import torch

    from omniisaacgymenvs_b200.robots.model import build_model
    from oracle import dynamics as dyn
    from oracle import locomotion_oracle as lo

    torch.set_num_threads(threads)
    spec = lo.SPECS[ROBOT]
    model = build_model(ROBOT)
    desc = dyn.make_desc(model)
    D = model.num_dof
    rng = np.random.default_rng(0)
    lim = model.dof_limits().astype(np.float64)
    pos = np.zeros((n_envs, 3)); pos[:, 2] = model.spawn_height
    quat = np.tile([[1.0, 0, 0, 0]], (n_envs, 1))
    vel = np.zeros((n_envs, 6))
    q = np.clip(rng.uniform(-0.2, 0.2, (n_envs, D)), lim[:, 0], lim[:, 1])
    qd = rng.uniform(-0.1, 0.1, (n_envs, D))
    sens = np.zeros((n_envs, model.num_sensors, 6))
    init = dict(pos=pos.copy(), quat=quat.copy(), q=q.copy())
    gears = torch.tensor(spec.joint_gears)
    s = {"potentials": torch.full((n_envs,), -1000.0 * 60.0), "progress_buf": torch.zeros(n_envs, dtype=torch.int64),
         "reset_buf": torch.zeros(n_envs, dtype=torch.int64),
         "targets": torch.tensor([1000.0, 0, 0]).repeat(n_envs, 1), "inv_start_rot": torch.tensor([1.0, 0, 0, 0]).repeat(n_envs, 1),
         "basis_vec0": torch.tensor([1.0, 0, 0]).repeat(n_envs, 1), "basis_vec1": torch.tensor([0.0, 0, 1]).repeat(n_envs, 1),
         "dof_lo": torch.from_numpy(lim[:, 0]).float(), "dof_hi": torch.from_numpy(lim[:, 1]).float()}

    def one_step():
        actions = torch.rand(n_envs, D) * 2 - 1
        ids = np.nonzero(s["reset_buf"].numpy())[0]
        if len(ids):
            pos[ids] = init["pos"][ids]; quat[ids] = init["quat"][ids]; vel[ids] = 0
            q[ids] = np.clip(init["q"][ids] + rng.uniform(-0.2, 0.2, (len(ids), D)), lim[:, 0], lim[:, 1])
            qd[ids] = rng.uniform(-0.1, 0.1, (len(ids), D))
            s["reset_buf"][ids] = 0; s["progress_buf"][ids] = 0
        eff = (actions * gears * spec.power_scale).double().numpy()
        dyn.step(desc, pos, quat, vel, q, qd, np.ascontiguousarray(eff), sens, 2, 0.0083, threads)
        s.update(pos=torch.from_numpy(pos).float(), quat=torch.from_numpy(quat).float(), vel=torch.from_numpy(vel).float(),
                 dof_pos=torch.from_numpy(q).float(), dof_vel=torch.from_numpy(qd).float(),
                 sensors=torch.from_numpy(sens).float(), actions=actions)
        out = lo.post_physics_step(spec, s)
        s["potentials"], s["progress_buf"], s["reset_buf"] = out["potentials"], out["progress_buf"], out["reset_buf"]
        return out

    for _ in range(warmup):
        one_step()
    t0 = time.perf_counter()
    for _ in range(steps):
        one_step()
    dt = time.perf_counter() - t0
    return n_envs * steps / dt, 1e3 * dt / steps


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    threads = os.cpu_count() or 1
    sample_envs = args.cpu_sample_envs
    rate, ms = cpu_env_step_rate(sample_envs, args.steps, max(args.warmup, 1), threads)
    line = {
        "impl": "reference", "metric": "humanoid_env_steps_per_sec", "value": rate, "unit": "env-steps/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"Humanoid full env step (K3 reset + 2x K2 articulation substeps + K1 obs/reward/reset), "
                               f"{args.envs_per_gpu} envs/GPU (BASELINE.json configs[2])",
                   "envs_per_gpu": args.envs_per_gpu, "sample_envs_per_step": sample_envs, "self_collision": True,
                   "arm": "CPU oracle port (float64 OpenMP dynamics incl. self-collision + torch obs/reward/reset), all host cores"},
        "cpu_baseline": {"value": rate, "unit": "env-steps/s", "cores": threads, "kind": "port",
                         "sample": f"{sample_envs} envs x {args.steps} full env steps (2 dynamics substeps + obs/reward/reset); "
                                   "PhysX/Isaac Sim not installable, float64 OpenMP oracle + torch CPU oracle stand in"},
        "e2e": {"value": rate, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
def k1_roofline(torch, device, n_envs, iters=20, warmup=5, flush=None):
    """Achieved bandwidth of K1 at n_envs via CUDA events on the launch stream (state + outputs > L2 at 1 Mi envs;
    at smaller counts `flush` -- a buffer larger than L2 -- is rewritten between launches, outside the event pairs)."""
    import ctypes as C

    from omniisaacgymenvs_b200 import _lib
    from omniisaacgymenvs_b200.utils.task_util import make_env

    import contextlib

    with contextlib.redirect_stdout(sys.stderr):
        env, task = make_env(ROBOT, num_envs=n_envs, device_id=device.index or 0)
    v = task.get_robot()
    g = torch.Generator(device=device).manual_seed(0)
    # synthetic state of SURVEY.md 8(d): everything random, heights straddling the termination height
    v.root_pos.normal_(generator=g); v.root_pos[:, 2].mul_(0.3).add_(1.0)
    v.root_quat.normal_(generator=g); v.root_quat.div_(v.root_quat.norm(dim=-1, keepdim=True))
    v.root_vel.normal_(generator=g); v.dof_vel.normal_(generator=g); v.sensor_forces.normal_(generator=g).mul_(10)
    v.dof_pos.uniform_(-1.0, 1.0, generator=g); task.actions.uniform_(-1.0, 1.0, generator=g)
    task.progress_buf.random_(0, 900, generator=g); task.reset_buf.zero_()
    lib = _lib.load()
    stream = _lib.current_stream_ptr(device)

    def launch():
        rc = lib.oige_obs_reward_reset_f32(C.byref(task._params), C.byref(v.state_struct), C.byref(task._buffers),
                                           n_envs, stream)
        _lib.check(rc, "K1")

    for _ in range(warmup):
        launch()
    torch.cuda.synchronize(device)
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in evs:
        if flush is not None:
            flush.zero_()
        a.record(); launch(); b.record()
    torch.cuda.synchronize(device)
    ms = [a.elapsed_time(b) for a, b in evs]
    del env, task, v
    torch.cuda.empty_cache()
    return statistics.mean(ms), min(ms)


def step_breakdown(torch, env, task, actions, iters=10):
    """Device time of each kernel of one step (hook-by-hook calls, CUDA events)."""
    import ctypes as C

    from omniisaacgymenvs_b200 import _lib

    lib = _lib.load()
    v = task.get_robot()
    dev = torch.device(task.device)
    stream = _lib.current_stream_ptr(dev)
    n = task.num_envs
    acc = {"reset_ms": 0.0, "articulation_ms": 0.0, "obs_reward_ms": 0.0}
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    for _ in range(iters):
        torch.cuda.synchronize(dev)
        ev[0].record()
        _lib.check(lib.oige_reset_masked_f32(C.byref(task._params), C.byref(v.state_struct), C.byref(task._buffers),
                                             C.byref(task._reset_init), None, None, 1, 1, 0, n, stream), "K3")
        ev[1].record()
        _lib.check(lib.oige_articulation_step_f32(v.model_handle, C.byref(task._params), C.byref(v.state_struct),
                                                  actions.data_ptr(), n, 2, env._world.get_physics_dt(), stream), "K2")
        ev[2].record()
        _lib.check(lib.oige_obs_reward_reset_f32(C.byref(task._params), C.byref(v.state_struct), C.byref(task._buffers),
                                                 n, stream), "K1")
        ev[3].record()
        torch.cuda.synchronize(dev)
        acc["reset_ms"] += ev[0].elapsed_time(ev[1]) / iters
        acc["articulation_ms"] += ev[1].elapsed_time(ev[2]) / iters
        acc["obs_reward_ms"] += ev[2].elapsed_time(ev[3]) / iters
    return acc


def timed_steps(torch, env, acts, flush, K, device):
    """K steps, per-step CUDA-event pairs, L2 flushed between steps outside the pairs; returns total device ms."""
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    for i in range(K):
        flush.zero_()
        a, b = evs[i]
        a.record()
        env.step(acts[i % len(acts)])
        b.record()
    torch.cuda.synchronize(device)
    return sum(a.elapsed_time(b) for a, b in evs)


def run_ours(args):
    import torch

    # host-buffer step knobs (read once when the library loads; the device-resident path does not use them): four chunks,
    # observations, rewards and reset flags stored straight into the pinned host buffers and actions read straight from
    # them -- the best of the combinations in profiles/r2_e2e_breakdown.txt.  Opt-in because the device copy of obs_buf is
    # then not written.
    os.environ.setdefault("OIGE_HOST_ZEROCOPY", "3")
    os.environ.setdefault("OIGE_HOST_CHUNKS", "4")
    from omniisaacgymenvs_b200 import _lib

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py (our arm) needs a CUDA device: there is no CPU fallback")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist_mod

        dist = dist_mod
        dist.init_process_group("nccl", device_id=device)

    from omniisaacgymenvs_b200.utils.task_util import make_env

    n = args.envs_per_gpu
    import contextlib

    with contextlib.redirect_stdout(sys.stderr):  # stdout carries exactly one JSON line
        env, task = make_env(ROBOT, num_envs=n, device_id=local_rank, seed=42, env_id_offset=rank * n,
                             global_num_envs=world * n)
        env.reset()
        if not args.no_cuda_graph:
            env.enable_cuda_graph()  # whole step (K3+clamp, K2, K1, hand-off copies) replayed as one graph
    K, W = args.steps, max(args.warmup, 3)
    D = task.num_actions
    gen = torch.Generator(device=device).manual_seed(1234 + rank)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)  # > 126 MB L2
    from omniisaacgymenvs_b200.utils.distributed import EpisodeStatsReducer, max_over_ranks

    reducer = EpisodeStatsReducer(device)

    def reduce_stats():
        reducer.reduce(task.read_episode_stats())

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(device)

    # ---- warm-up (the clock sampler starts here so that short timed regions still see samples under load) ----
    clocks = ClockSampler(local_rank)
    clocks.__enter__()
    for _ in range(W):
        env.step(torch.rand(n, D, device=device, generator=gen) * 2 - 1)
    reduce_stats()  # first NCCL collective of the process (lazy connection set-up) happens here, not in the timed region
    if reducer._side is not None:
        reducer._side.synchronize()
    barrier()

    # ---- timed region: device-resident inputs, per-step event pairs, L2 flushed between steps ----
    acts = [torch.rand(n, D, device=device, generator=gen) * 2 - 1 for _ in range(min(K, 64))]

    def protocol(k_steps):
        evs_ = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(k_steps)]
        barrier()
        lo = clocks.mark()
        t0 = time.perf_counter()
        for i in range(k_steps):
            flush.zero_()
            a, b = evs_[i]
            a.record()
            env.step(acts[i % len(acts)])
            b.record()
            if (i + 1) % 16 == 0:
                reduce_stats()
        barrier()
        return evs_, lo, clocks.mark(), time.perf_counter() - t0

    # Untimed rehearsal of the same protocol (more warm-up, on top of the W steps above): an event pair brackets
    # "host enqueues the step" as well as the step, so in the first process on a fresh box -- host code paths still
    # paging in -- the host falls behind the device and the pairs measure launch latency (seen once in round 3:
    # 0.33 ms per step in the first run on a box, 0.10 ms in the second).  The rehearsal pays those first-touch costs.
    protocol(max(K, 16))
    launches0 = _lib.launch_count()
    replays0 = getattr(env, "graph_replays", 0)
    evs, clk_lo, clk_hi, t_wall = protocol(K)
    launches = _lib.launch_count() - launches0
    if getattr(env, "_graph", None) is not None:  # kernels inside graph replays are not host-side launches
        launches += K * env.graph_kernels
    dev_ms = sum(a.elapsed_time(b) for a, b in evs)
    dev_ms = max_over_ranks(dev_ms, device)
    ms_per_step = dev_ms / K
    value = world * n * K / (dev_ms / 1e3)
    reduce_stats()
    episode_stats = reducer.result()

    # ---- end-to-end: HOST buffers through the C-ABI host entry point, copies inside the timed region ----
    a_host = [(torch.rand(n, D) * 2 - 1).pin_memory() for _ in range(4)]
    obs_h = torch.empty(n, task.num_observations).pin_memory()
    rew_h = torch.empty(n).pin_memory()
    rst_h = torch.empty(n, dtype=torch.int64).pin_memory()
    env.disable_cuda_graph()
    if not args.no_cuda_graph:  # one captured graph per ring slot: H2D, 4 chunk streams of kernels, D2H
        for a in a_host:
            env.enable_cuda_graph_host(a, obs_h, rew_h, rst_h)
    for i in range(3):
        env.step_host(a_host[i % 4], obs_h, rew_h, rst_h)
    clocks.interval = 0.1  # one NVML polling thread per rank: keep it off the host cores while the copies are timed
    barrier()
    t0 = time.perf_counter()
    for i in range(K):
        env.step_host(a_host[i % 4], obs_h, rew_h, rst_h)
        torch.cuda.synchronize(device)  # the caller reads obs/rew/reset every step
    e2e_s = time.perf_counter() - t0  # this rank's own loop; the slowest rank sets the job's rate (max over ranks)
    e2e_value = world * n * K / max_over_ranks(e2e_s, device)
    clocks.interval = 0.001
    h2d = n * D * 4
    d2h = n * task.num_observations * 4 + n * 4 + n * 8

    # ---- sustained: the same CUDA-graph step looped for >= 2 s (hot L2, no flush), with its own clock record ----
    env.disable_cuda_graph_host()
    if not args.no_cuda_graph:
        env.enable_cuda_graph()
    for _ in range(W):
        env.step(acts[0])
    barrier()
    sus_lo = clocks.mark()
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sus_steps, t0 = 0, time.perf_counter()
    s0.record()
    while time.perf_counter() - t0 < args.sustained_s:
        for i in range(256):
            env.step(acts[i % len(acts)])
        sus_steps += 256
        torch.cuda.synchronize(device)  # bounds the launch queue; ~25 ms of work per sync
    s1.record()
    torch.cuda.synchronize(device)
    sus_ms = max_over_ranks(s0.elapsed_time(s1), device)
    sus_hi = clocks.mark()
    sustained = {"value": world * n * sus_steps / (sus_ms / 1e3), "unit": "env-steps/s", "steps": sus_steps,
                 "seconds": sus_ms / 1e3, "us_per_step": 1e3 * sus_ms / max(sus_steps, 1), "l2": "hot (no flush)",
                 "clocks": clocks.summary(sus_lo, sus_hi)}
    clocks.__exit__()

    line = None
    if rank == 0:
        breakdown = step_breakdown(torch, env, task, acts[0])
        del env, task
        torch.cuda.empty_cache()
        # `value`'s twin without self-collision (the mechanical system of round 1)
        nosc = None
        if not args.no_extras:
            with contextlib.redirect_stdout(sys.stderr):
                env2, task2 = make_env(ROBOT, num_envs=n, device_id=local_rank, seed=42, env_id_offset=rank * n,
                                       global_num_envs=world * n,
                                       overrides={"task.sim.Humanoid.enable_self_collisions": False})
                env2.reset()
                if not args.no_cuda_graph:
                    env2.enable_cuda_graph()
            for _ in range(W):
                env2.step(acts[0])
            torch.cuda.synchronize(device)
            ms2 = timed_steps(torch, env2, acts, flush, 20, device)
            nosc = {"value": n * 20 / (ms2 / 1e3), "unit": "env-steps/s (this rank)", "ms_per_step": ms2 / 20, "steps": 20}
            del env2, task2
            torch.cuda.empty_cache()
        peak, peak_src = read_peaks()
        ncu = read_k2f_ncu()
        step_s = ms_per_step * 1e-3
        sm_mhz = clocks.summary(clk_lo, clk_hi)["sm_mhz"] or clocks.sm_max or 1965.0
        n_sm = torch.cuda.get_device_properties(device).multi_processor_count
        issue_peak = n_sm * 4 * sm_mhz * 1e6  # one warp-instruction per scheduler per cycle, 4 schedulers per SM
        scale = n / 16384.0  # the capture is of a 16384-env launch
        achieved = STEP_BYTES[ROBOT] * n / step_s / 1e9
        roof = {"bound": "hbm", "kernel": "env_step_role_kernel<6,3,false,128,1,true,true> (K2-F: the whole env step in one launch)",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": ncu["dram_bytes"] * scale, "traffic_source": ncu["source"], "peak_source": peak_src,
                "envs": n, "bytes_per_env": STEP_BYTES[ROBOT], "avg_ms": ms_per_step,
                "limiter": "issue / dependency latency, not HBM: see `issue`",
                "note": "algorithmic bytes = state in + state, observations, reward, flags out (944 B per env-step) over the "
                        "device time of the step; the kernel keeps everything else on chip (ncu DRAM traffic ~ the "
                        "algorithmic bytes, no spill traffic)"}
        issue = {"warp_inst_per_launch": ncu["warp_inst"] * scale, "source": ncu["source"],
                 "achieved_ginst_s": ncu["warp_inst"] * scale / step_s / 1e9, "peak_ginst_s": issue_peak / 1e9,
                 "frac": ncu["warp_inst"] * scale / step_s / issue_peak, "ncu_issue_util_pct": ncu["issue_util_pct"],
                 "ncu_time_us": ncu["time_us"], "sm_mhz": sm_mhz,
                 "note": "warp-instructions of the committed ncu capture / live device time of the step, against "
                         "SMs x 4 schedulers x SM clock; the step time includes the graph's hand-off copies"}
        roof_k1 = None
        if not args.no_extras:
            k1_ms, k1_best = k1_roofline(torch, device, ROOFLINE_ENVS)
            a1 = K1_BYTES[ROBOT] * ROOFLINE_ENVS / (k1_ms * 1e-3) / 1e9
            k1s_ms, k1s_best = k1_roofline(torch, device, n, flush=flush)
            a2 = K1_BYTES[ROBOT] * n / (k1s_ms * 1e-3) / 1e9
            roof_k1 = {"bound": "hbm", "kernel": "obs_reward_reset_kernel<21,2,32> (stand-alone obs/reward/reset pass; inside "
                                                  "the fused step its arithmetic runs in K2-F's epilogue)",
                       "achieved": a1, "peak": peak, "unit": "GB/s", "frac": a1 / peak, "envs": ROOFLINE_ENVS,
                       "bytes_per_env": K1_BYTES[ROBOT], "avg_ms": k1_ms, "best_ms": k1_best,
                       "traffic": K1_TRAFFIC_BYTES_1M, "traffic_source": K1_TRAFFIC_SOURCE,
                       "at_bench_envs": {"envs": n, "achieved": a2, "frac": a2 / peak, "avg_ms": k1s_ms, "best_ms": k1s_best,
                                         "l2": "flushed between launches"}}
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            rate, cms = cpu_env_step_rate(args.cpu_sample_envs, 30, 2, threads)
            rate1, _ = cpu_env_step_rate(2048, 5, 1, 1)
            cpu = {"value": rate, "unit": "env-steps/s", "cores": threads, "kind": "port",
                   "value_1thread": rate1, "sample_1thread": "2048 envs x 5 full env steps, 1 thread",
                   "sample": f"{args.cpu_sample_envs} envs x 30 full env steps on the host cores "
                             "(float64 OpenMP dynamics oracle incl. self-collision, 2 substeps, + torch CPU obs/reward/reset oracle)"}
        line = {
            "metric": "humanoid_env_steps_per_sec", "value": value, "unit": "env-steps/s", "n_gpus": world,
            "steps": K, "warmup": W, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"Humanoid full env step (K3 reset + 2x K2 articulation substeps + K1 obs/reward/reset), "
                                   f"{n} envs/GPU (BASELINE.json configs[2])",
                       "envs_per_gpu": n, "global_envs": world * n, "parallelism": f"env-index sharding x{world}",
                       "kernels": "one launch per step: K2-F env_step_role_kernel (reset + clamp, 2 substeps, obs/reward/done fused)",
                       "self_collision": True,
                       "l2": "flushed between steps (256 MiB memset outside the per-step CUDA-event pairs)",
                       "actions": "uniform random policy, device-resident",
                       "cuda_graph": not args.no_cuda_graph,
                       "physics_parity": "vs PhysX UNPINNED: PhysX 5 is closed and absent; the articulation step is this repo's "
                                         "own solver, checked against its float64 oracle and model-level invariants (DESIGN.md section 4)"},
            "clocks": clocks.summary(clk_lo, clk_hi),
            "e2e": {"value": e2e_value, "unit": "env-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "timing": "per-rank wall clock around K host-buffer steps (one stream sync per step), max over ranks",
                    "knobs": {k: os.environ.get(k) for k in ("OIGE_HOST_ZEROCOPY", "OIGE_HOST_CHUNKS")}},
            "gpu_launches": int(launches),
            "roofline": roof, "issue": issue, "roofline_k1": roof_k1, "sustained": sustained, "no_self_collision": nosc,
            "cpu_baseline": cpu, "step_breakdown": breakdown,
            "wall_s_timed_region": t_wall,
            "episode_stats": episode_stats,
        }
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if line is not None:
        print(json.dumps(line), flush=True)
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--envs-per-gpu", type=int, default=ENVS_PER_GPU)
    ap.add_argument("--cpu-sample-envs", type=int, default=ENVS_PER_GPU)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-cuda-graph", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the K1 roofline and no-self-collision side measurements")
    ap.add_argument("--sustained-s", type=float, default=2.0, help="length of the sustained (loaded-GPU) loop")
    args = ap.parse_args()
    # stdout carries exactly ONE JSON line: libraries that write to the C-level stdout (NCCL prints its version
    # banner there) are sent to stderr by pointing fd 1 at fd 2 and giving Python's sys.stdout the original fd
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = real_stdout
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
