import contextlib, io, os, sys
sys.path.insert(0, "/root/repo")
import torch
from omniisaacgymenvs_b200.utils.task_util import make_env
n=16384
with contextlib.redirect_stdout(io.StringIO()):
    env, task = make_env("Humanoid", num_envs=n); env.reset()
a=(torch.rand(n,21)*2-1).pin_memory(); o=torch.empty(n,87).pin_memory(); r=torch.empty(n).pin_memory(); d=torch.empty(n,dtype=torch.int64).pin_memory()
for i in range(8):
    env.step_host(a,o,r,d); torch.cuda.synchronize()
# the same step replayed from a CUDA graph (no CPU issue cost): timeline of the last replay
from omniisaacgymenvs_b200 import _lib
env.enable_cuda_graph_host(a, o, r, d)
for i in range(6):
    env.step_host(a, o, r, d); torch.cuda.synchronize()
print("graph replay:", file=sys.stderr, flush=True)
rc = _lib.load().oige_host_trace_dump()
if rc != 0:
    print("  (no timeline: events recorded by a graph replay cannot be timed on this driver, rc=%d)" % rc, file=sys.stderr)
