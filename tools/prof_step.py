"""Short profiling target (run under ncu on the GPU box): a few full env steps at 16384 envs and a
few K1 launches at 1,048,576 envs.  Usage: python tools/prof_step.py [steps] [k1_launches]"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from omniisaacgymenvs_b200 import _lib
from omniisaacgymenvs_b200.utils.task_util import make_env

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 4
k1n = int(sys.argv[2]) if len(sys.argv) > 2 else 3
env, task = make_env("Humanoid", num_envs=16384)
env.reset()
for _ in range(steps):
    env.step(torch.rand(16384, 21, device=task.device) * 2 - 1)
torch.cuda.synchronize()
if k1n:
    del env, task
    n = 1 << 20
    env, task = make_env("Humanoid", num_envs=n)
    v = task.get_robot()
    v.root_pos.normal_(); v.root_pos[:, 2].mul_(0.3).add_(1.0)
    v.root_quat.normal_(); v.root_quat.div_(v.root_quat.norm(dim=-1, keepdim=True))
    v.root_vel.normal_(); v.dof_vel.normal_(); v.dof_pos.uniform_(-1, 1); task.actions.uniform_(-1, 1)
    lib = _lib.load()
    for _ in range(k1n):
        _lib.check(lib.oige_obs_reward_reset_f32(C.byref(task._params), C.byref(v.state_struct), C.byref(task._buffers),
                                                 n, _lib.current_stream_ptr(torch.device(task.device))), "K1")
    torch.cuda.synchronize()
print("ok")
