"""K2 timing (GPU box): device time of the articulation step at 16384 envs (2 substeps)."""
import ctypes as C, os, sys, statistics, io, contextlib
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from omniisaacgymenvs_b200 import _lib
from omniisaacgymenvs_b200.utils.task_util import make_env
lib = _lib.load()
for robot in (sys.argv[1].split(",") if len(sys.argv) > 1 else ["Humanoid", "Ant"]):
    for n in [int(x) for x in os.environ.get("K2_NS", "16384,65536").split(",")]:
        with contextlib.redirect_stdout(io.StringIO()):
            env, task = make_env(robot, num_envs=n)
            env.reset()
        dev = torch.device(task.device); v = task.get_robot()
        for _ in range(30):  # reach a typical mixed state (some fallen, some standing)
            env.step(torch.rand(n, task.num_actions, device=dev) * 2 - 1)
        a = torch.rand(n, task.num_actions, device=dev) * 2 - 1
        stream = _lib.current_stream_ptr(dev); ms = []
        for it in range(14):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            _lib.check(lib.oige_articulation_step_f32(v.model_handle, C.byref(task._params), C.byref(v.state_struct), a.data_ptr(), n, 2, 0.0083, stream), "k2")
            e1.record(); torch.cuda.synchronize()
            if it >= 4: ms.append(e0.elapsed_time(e1))
        t = statistics.mean(ms)
        print(f"{robot} n={n}: K2 {t*1e3:.1f} us  ({n/t/1e3:.1f} M env-steps/s, {2*n/t/1e3:.1f} M substeps/s)", flush=True)
        del env, task, v; torch.cuda.empty_cache()
