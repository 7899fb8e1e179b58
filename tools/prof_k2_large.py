"""Profiling target: K2 at a large env count (default 262144 Humanoids), a few launches after a short rollout."""
import ctypes as C, os, sys, io, contextlib
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from omniisaacgymenvs_b200 import _lib
from omniisaacgymenvs_b200.utils.task_util import make_env
n = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
with contextlib.redirect_stdout(io.StringIO()):
    env, task = make_env("Humanoid", num_envs=n); env.reset()
for _ in range(10):
    env.step(torch.rand(n, 21, device=task.device) * 2 - 1)
v = task.get_robot(); a = torch.rand(n, 21, device=task.device) * 2 - 1
lib = _lib.load()
for _ in range(3):
    _lib.check(lib.oige_articulation_step_f32(v.model_handle, C.byref(task._params), C.byref(v.state_struct), a.data_ptr(), n, 2, 0.0083,
                                              _lib.current_stream_ptr(torch.device(task.device))), "k2")
torch.cuda.synchronize(); print("ok")
