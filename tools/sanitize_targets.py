"""Small launches of every kernel family for compute-sanitizer (racecheck / synccheck / memcheck): ragged env counts so
that both the bulk-copy (TMA) tiles and the plain tail path run.  argv[1]: which = k2f | k2modes | passes | all.
Usage (one tool per gpurun call):  compute-sanitizer --tool racecheck python tools/sanitize_targets.py k2f"""
import ctypes as C, contextlib, io, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from omniisaacgymenvs_b200 import _lib
from omniisaacgymenvs_b200.utils.task_util import make_env

which = sys.argv[1] if len(sys.argv) > 1 else "all"
lib = _lib.load()


def steps(robot, n, k=3, **kw):
    with contextlib.redirect_stdout(io.StringIO()):
        env, task = make_env(robot, num_envs=n, **kw)
        env.reset()
    dev = torch.device(task.device)
    for _ in range(k):
        env.step(torch.rand(n, task.num_actions, device=dev) * 2 - 1)
    torch.cuda.synchronize()
    return env, task


if which in ("k2f", "all"):
    # whole-step kernel: 128-env CTAs need > 3 x SMs groups; 32-env CTAs otherwise; ragged tail (non-bulk path)
    for robot, n in (("Humanoid", 4096 + 37), ("Ant", 1024 + 5), ("Anymal", 1024 + 5)):
        steps(robot, n)
        print("k2f", robot, n, "ok", flush=True)
    os.environ["OIGE_K2_EPC"] = "128"; lib.oige_reload_env()
    steps("Humanoid", 4096 + 37)
    os.environ.pop("OIGE_K2_EPC"); lib.oige_reload_env()
    print("k2f Humanoid 128-env CTAs ok", flush=True)
if which in ("k2modes", "all"):
    for mode in ("warp", "thread"):
        os.environ["OIGE_K2_MODE"] = mode; lib.oige_reload_env()
        for robot, n in (("Humanoid", 256 + 5), ("Ant", 256 + 5)):
            steps(robot, n)
            print("k2", mode, robot, n, "ok", flush=True)
    os.environ.pop("OIGE_K2_MODE"); lib.oige_reload_env()
if which in ("passes", "all"):
    # K3 + K1 + DR + Cartpole / Anymal passes (K2-W below the role kernel's threshold)
    steps("Humanoid", 512 + 3)
    steps("Cartpole", 512 + 3)
    steps("Anymal", 512 + 3)
    print("passes ok", flush=True)
print("done", _lib.nonfinite_env_count())
