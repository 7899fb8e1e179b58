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
