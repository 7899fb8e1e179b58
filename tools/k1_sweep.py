"""K1 bandwidth sweep (GPU box): achieved GB/s per tile size and env count."""
import ctypes as C, os, sys, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from omniisaacgymenvs_b200 import _lib
from omniisaacgymenvs_b200.utils.task_util import make_env
import io, contextlib
lib = _lib.load()
robots = sys.argv[1].split(",") if len(sys.argv) > 1 else ["Humanoid"]
for robot, bpe in (("Humanoid", 772), ("Ant", 556)):
    if robot not in robots: continue
    for n in (4096, 16384, 65536, 262144, 1 << 20, 1 << 21):
        with contextlib.redirect_stdout(io.StringIO()):
            env, task = make_env(robot, num_envs=n)
        v = task.get_robot(); dev = torch.device(task.device)
        v.root_pos.normal_(); v.root_pos[:, 2].mul_(0.3).add_(1.0)
        v.root_quat.normal_(); v.root_quat.div_(v.root_quat.norm(dim=-1, keepdim=True))
        v.root_vel.normal_(); v.dof_vel.normal_(); v.dof_pos.uniform_(-1, 1); task.actions.uniform_(-1, 1)
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
        stream = _lib.current_stream_ptr(dev)
        out = [f"{robot} n={n:8d}"]
        for tile in (32, 64):
            ms = []
            for it in range(12):
                if n * bpe < (200 << 20): flush.zero_()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                _lib.check(lib.oige_obs_reward_reset_tile_f32(C.byref(task._params), C.byref(v.state_struct), C.byref(task._buffers), n, tile, stream), "k1")
                b.record(); torch.cuda.synchronize()
                if it >= 2: ms.append(a.elapsed_time(b))
            t = statistics.mean(ms)
            out.append(f"tile{tile}: {t*1e3:8.1f} us {n*bpe/t/1e6:7.0f} GB/s")
        print(" | ".join(out), flush=True)
        del env, task, v, flush; torch.cuda.empty_cache()
