import sys, io, contextlib; import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from omniisaacgymenvs_b200.utils.task_util import make_env
for ov in (None, {"task.sim.Humanoid.enable_self_collisions": False}):
    print("=== overrides", ov, file=sys.stderr)
    with contextlib.redirect_stdout(io.StringIO()):
        env, task = make_env("Humanoid", num_envs=64, overrides=ov)
