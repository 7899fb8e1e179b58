"""Where the end-to-end (host-buffer) step goes (GPU box): copies alone, device step alone, step_host eager / graph."""
import contextlib, io, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from omniisaacgymenvs_b200.utils.task_util import make_env

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
dev = torch.device("cuda:0")
with contextlib.redirect_stdout(io.StringIO()):
    env, task = make_env("Humanoid", num_envs=n); env.reset()
A, O = task.num_actions, task.num_observations
ring = [(torch.rand(n, A) * 2 - 1).pin_memory() for _ in range(4)]
obs_h = torch.empty(n, O).pin_memory(); rew_h = torch.empty(n).pin_memory(); rst_h = torch.empty(n, dtype=torch.int64).pin_memory()
a_d = torch.empty(n, A, device=dev)


def wall(fn, iters=200):
    for _ in range(10): fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(iters):
        fn(i); torch.cuda.synchronize()
    return (time.perf_counter() - t0) / iters * 1e6


def h2d(i=0): a_d.copy_(ring[i % 4], non_blocking=True)
def d2h(i=0):
    obs_h.copy_(task.obs_buf, non_blocking=True); rew_h.copy_(task.rew_buf, non_blocking=True); rst_h.copy_(task.reset_buf, non_blocking=True)
def dev_step(i=0): env.step(a_d)
def host_step(i=0): env.step_host(ring[i % 4], obs_h, rew_h, rst_h)
def serial(i=0): h2d(i); env.step(a_d); d2h(i)

print(f"Humanoid n={n}; wall us per call incl. one synchronize")
print(f"  sync only                 {wall(lambda i=0: None):8.1f}")
print(f"  H2D actions ({n*A*4/1e6:.2f} MB)      {wall(h2d):8.1f}")
print(f"  D2H obs+rew+rst ({(n*O*4+n*12)/1e6:.2f} MB) {wall(d2h):8.1f}")
print(f"  device step, eager        {wall(dev_step):8.1f}")
print(f"  H2D + step + D2H, eager   {wall(serial):8.1f}")
print(f"  step_host eager (4 chunks){wall(host_step):8.1f}")
for a in ring: env.enable_cuda_graph_host(a, obs_h, rew_h, rst_h)
print(f"  step_host CUDA graph      {wall(host_step):8.1f}")
env.disable_cuda_graph_host(); env.enable_cuda_graph()
print(f"  device step, CUDA graph   {wall(dev_step):8.1f}")
print(f"  H2D + graph step + D2H    {wall(serial):8.1f}")
