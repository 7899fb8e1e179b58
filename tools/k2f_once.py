"""A few whole-step launches (K2-F, env_step_role_kernel) in a typical mixed state: target of an
`ncu -k regex:env_step_role --launch-skip 20 -c 1` capture.  argv: robot, envs, [physics]: `physics` launches the
kernel through oige_articulation_step_f32 (prologue / epilogue switched off) instead of the whole step."""
import ctypes as C, os, sys, io, contextlib
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from omniisaacgymenvs_b200 import _lib
from omniisaacgymenvs_b200.utils.task_util import make_env
lib = _lib.load()
robot = sys.argv[1] if len(sys.argv) > 1 else "Humanoid"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
physics = "physics" in sys.argv[3:]
nosc = "nosc" in sys.argv[3:]  # Humanoid with enable_self_collisions off
with contextlib.redirect_stdout(io.StringIO()):
    env, task = make_env(robot, num_envs=n, overrides={"task.sim.Humanoid.enable_self_collisions": False} if nosc else None)
    env.reset()
dev = torch.device(task.device); v = task.get_robot()
for _ in range(20):
    env.step(torch.rand(n, task.num_actions, device=dev) * 2 - 1)
a = torch.rand(n, task.num_actions, device=dev) * 2 - 1
for _ in range(3):
    if physics:
        _lib.check(lib.oige_articulation_step_f32(v.model_handle, C.byref(task._params), C.byref(v.state_struct), a.data_ptr(), n, 2,
                                                  0.0083, _lib.current_stream_ptr(dev)), "k2")
    else:
        env.step(a)
torch.cuda.synchronize()
print("ok", _lib.nonfinite_env_count())
