"""Random-policy loop (reference: scripts/random_policy.py:39-68) on the B200 env step.

    python -m omniisaacgymenvs_b200.scripts.random_policy task=Humanoid num_envs=4096 steps=200

Same loop body as the reference (pre_physics_step -> world.step -> post_physics_step through the task
hooks), except that it honours controlFrequencyInv and stops after `steps` iterations because there is
no Kit window to close.  Overrides use the reference's `key=value` CLI form.
"""
import sys
import time

import numpy as np
import torch

from omniisaacgymenvs_b200.envs.vec_env_rlgames import VecEnvRLGames
from omniisaacgymenvs_b200.utils.config import load_config
from omniisaacgymenvs_b200.utils.task_util import initialize_task


def parse_overrides(argv):
    out = {}
    for a in argv:
        k, _, v = a.partition("=")
        try:
            out[k] = int(v)
        except ValueError:
            try:
                out[k] = float(v)
            except ValueError:
                out[k] = {"True": True, "False": False}.get(v, v)
    return out


def main(argv=None):
    ov = parse_overrides(sys.argv[1:] if argv is None else argv)
    task_name = ov.pop("task", "Humanoid")
    steps = int(ov.pop("steps", 100))
    cfg_dict = load_config(task_name, ov)
    env = VecEnvRLGames(headless=cfg_dict["headless"], sim_device=cfg_dict["device_id"])
    task = initialize_task(cfg_dict, env)
    t0 = time.perf_counter()
    it = 0
    while env._simulation_app.is_running() and it < steps:
        if env._world.is_playing():
            if env._world.current_time_step_index == 0:
                env._world.reset(soft=True)
            actions = torch.tensor(np.array([env.action_space.sample() for _ in range(env.num_envs)]),
                                   device=task.rl_device)
            env._task.pre_physics_step(actions)
            for _ in range(task.control_frequency_inv):
                env._world.step(render=False)
                env.sim_frame_count += 1
            env._task.post_physics_step()
        it += 1
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"{it} steps of {env.num_envs} envs in {dt:.2f} s; mean reward {float(task.rew_buf.mean()):.3f}; "
          f"resets pending {int(task.reset_buf.sum())}")
    env._simulation_app.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
