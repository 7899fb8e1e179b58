"""Why does the host-buffer step slow down when several ranks run at once?  (VERDICT r1 "What's weak" 6)

    python tools/e2e_scaling.py                                             # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 --master-port 29533 \
        tools/e2e_scaling.py

Every rank, all ranks AT THE SAME TIME (barrier before each section), on its own GPU with its own pinned buffers:
  1. H2D of the bench's action block (1.38 MB) and D2H of its result block (5.9 MB) and of 64 MiB: single-copy latency
     and back-to-back throughput (CUDA events) -- the concurrent-copy ceiling of the box;
  2. the same with the OTHER ranks idle (rank r measures alone in turn) -- the solo figure on the same box;
  3. the host-buffer env step (`VecEnvRLGames.step_host`, CUDA-graph replay + one stream sync per step), per rank,
     concurrently and solo; wall clock per step.
Rank 0 prints min / median / max over ranks.  Also prints each GPU's CPU affinity / NUMA node (NVML) and the process's own
affinity, so that PCIe uplink sharing, host memory bandwidth and host cores can be told apart."""
import contextlib
import io
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from omniisaacgymenvs_b200.utils.task_util import make_env

rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist = None
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=dev)


def barrier():
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize(dev)


def gather(x):
    t = torch.tensor([float(x)], device=dev, dtype=torch.float64)
    if dist is None:
        return [float(x)]
    out = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(out, t)
    return [float(o) for o in out]


def report(name, vals, unit, fmt="{:8.1f}"):
    if rank == 0:
        v = sorted(vals)
        print(f"  {name:58s} min {fmt.format(v[0])}  median {fmt.format(statistics.median(v))}  max {fmt.format(v[-1])} {unit}", flush=True)


def copy_rates(size, h2d):
    h = torch.empty(size, dtype=torch.uint8).pin_memory(); d = torch.empty(size, dtype=torch.uint8, device=dev)
    dst, src = (d, h) if h2d else (h, d)
    for _ in range(3):
        dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize(dev)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); dst.copy_(src, non_blocking=True); b.record(); torch.cuda.synchronize(dev)
    one = a.elapsed_time(b)
    reps = 20
    a.record()
    for _ in range(reps):
        dst.copy_(src, non_blocking=True)
    b.record(); torch.cuda.synchronize(dev)
    return one * 1e3, size / (a.elapsed_time(b) / reps) / 1e6  # us, GB/s


def section(concurrent, fn):
    """fn() on every rank at once, or on one rank at a time while the others wait at the barrier."""
    if concurrent or world == 1:
        barrier(); r = fn(); barrier()
        return r
    res = None
    for turn in range(world):
        barrier()
        if turn == rank:
            res = fn()
    barrier()
    return res


if rank == 0:
    print(f"# {torch.cuda.get_device_name(local)} x{world}, host cores {os.cpu_count()}, this process may run on {len(os.sched_getaffinity(0))} of them")
try:
    import pynvml
    pynvml.nvmlInit()
    h = pynvml.nvmlDeviceGetHandleByIndex(local)
    aff = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
    cores = [i for i in range(os.cpu_count()) if (aff[i // 64] >> (i % 64)) & 1]
    try:
        numa = pynvml.nvmlDeviceGetNumaNodeId(h)
    except Exception:
        numa = "?"
    pci = pynvml.nvmlDeviceGetPciInfo(h).busId
    line = f"gpu {local}: pci {pci if isinstance(pci, str) else pci.decode()}, numa node {numa}, cpu affinity {cores[0]}-{cores[-1]} ({len(cores)} cores)"
except Exception as e:  # noqa
    line = f"gpu {local}: nvml unavailable ({e})"
for r in range(world):
    barrier()
    if r == rank:
        print("# " + line, flush=True)
barrier()

for mode, conc in (("CONCURRENT (all ranks at once)", True), ("SOLO (one rank at a time)", False)):
    if world == 1 and not conc:
        break
    if rank == 0:
        print(f"== copies, {mode}")
    for size, h2d, name in ((1376256, True, "H2D 1.38 MB (action block of 16384 Humanoids)"), (5898240, False, "D2H 5.90 MB (obs + rew + resets)"),
                            (64 << 20, True, "H2D 64 MiB"), (64 << 20, False, "D2H 64 MiB")):
        one, gbs = section(conc, lambda: copy_rates(size, h2d))
        report(name + ": single copy", gather(one), "us")
        report(name + ": back-to-back", gather(gbs), "GB/s", "{:8.2f}")

n = 16384
with contextlib.redirect_stdout(io.StringIO()):
    env, task = make_env("Humanoid", num_envs=n, device_id=local, seed=42, env_id_offset=rank * n, global_num_envs=world * n)
    env.reset()
D = task.num_actions
a_host = [(torch.rand(n, D) * 2 - 1).pin_memory() for _ in range(4)]
obs_h = torch.empty(n, task.num_observations).pin_memory(); rew_h = torch.empty(n).pin_memory()
rst_h = torch.empty(n, dtype=torch.int64).pin_memory()
for a in a_host:
    env.enable_cuda_graph_host(a, obs_h, rew_h, rst_h)


def host_steps(k=200):
    for i in range(5):
        env.step_host(a_host[i % 4], obs_h, rew_h, rst_h)
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    for i in range(k):
        env.step_host(a_host[i % 4], obs_h, rew_h, rst_h)
        torch.cuda.synchronize(dev)
    return (time.perf_counter() - t0) / k * 1e6


def device_steps(k=200):
    a = torch.rand(n, D, device=dev) * 2 - 1
    for _ in range(5):
        env.step(a)
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(k):
        env.step(a)
    e1.record(); torch.cuda.synchronize(dev)
    return e0.elapsed_time(e1) / k * 1e3


for mode, conc in (("CONCURRENT", True), ("SOLO", False)):
    if world == 1 and not conc:
        break
    us = section(conc, host_steps)
    if rank == 0:
        print(f"== host-buffer env step (16384 Humanoids per rank), {mode}")
    report("wall clock per step_host + sync", gather(us), "us")
    report("=> env-steps/s per rank", [n / u for u in gather(us)], "M", "{:8.2f}")
env.disable_cuda_graph_host()
env.enable_cuda_graph()
us = section(True, device_steps)
if rank == 0:
    print("== device-resident step (CUDA graph, hot L2), CONCURRENT")
report("device time per step", gather(us), "us")
if dist is not None:
    dist.destroy_process_group()
