"""Training entry (reference: scripts/rlgames_train.py:47-134) with the B200 env step.

    python -m omniisaacgymenvs_b200.scripts.rlgames_train task=Humanoid num_envs=16384 train_yaml=/path/HumanoidPPO.yaml

Builds the env exactly like the reference (VecEnvRLGames -> initialize_task -> set_seed), registers it
with rl_games under 'rlgpu' and runs the Runner.  rl_games itself is an external dependency of the
reference (setup.py:17) and is not vendored: without it this script stops with a clear message after
the environment has been built and smoke-stepped.
"""
import sys

import torch
import yaml

from omniisaacgymenvs_b200.envs.vec_env_rlgames import VecEnvRLGames
from omniisaacgymenvs_b200.scripts.random_policy import parse_overrides
from omniisaacgymenvs_b200.utils.config import load_config
from omniisaacgymenvs_b200.utils.maths import set_seed
from omniisaacgymenvs_b200.utils.rlgames.rlgames_utils import HAVE_RL_GAMES, RLGPUAlgoObserver, RLGPUEnv
from omniisaacgymenvs_b200.utils.task_util import initialize_task


class RLGTrainer:
    def __init__(self, cfg_dict, train_cfg):
        self.cfg_dict, self.rlg_config_dict = cfg_dict, train_cfg

    def launch_rlg(self, env):
        from rl_games.common import env_configurations, vecenv

        vecenv.register("RLGPU", lambda config_name, num_actors, **kwargs: RLGPUEnv(config_name, num_actors, **kwargs))
        env_configurations.register("rlgpu", {"vecenv_type": "RLGPU", "env_creator": lambda **kwargs: env})

    def run(self):
        from rl_games.torch_runner import Runner

        runner = Runner(RLGPUAlgoObserver())
        runner.load(self.rlg_config_dict)
        runner.reset()
        runner.run({"train": not self.cfg_dict["test"], "play": self.cfg_dict["test"],
                    "checkpoint": self.cfg_dict["checkpoint"], "sigma": None})


def main(argv=None):
    ov = parse_overrides(sys.argv[1:] if argv is None else argv)
    task_name = ov.pop("task", "Humanoid")
    train_yaml = ov.pop("train_yaml", None)
    cfg_dict = load_config(task_name, ov)
    env = VecEnvRLGames(headless=cfg_dict["headless"], sim_device=cfg_dict["device_id"])
    task = initialize_task(cfg_dict, env)
    cfg_dict["seed"] = set_seed(cfg_dict["seed"], torch_deterministic=cfg_dict["torch_deterministic"])
    if not HAVE_RL_GAMES or train_yaml is None:
        obs = env.reset()
        env.step(torch.zeros(env.num_envs, task.num_actions, device=task.device))
        print(f"environment ready: obs {tuple(obs['obs'].shape)}; rl_games "
              f"{'missing' if not HAVE_RL_GAMES else 'present'} / train_yaml {train_yaml}: nothing to train with")
        return 2
    with open(train_yaml) as f:
        train_cfg = yaml.safe_load(f)
    trainer = RLGTrainer(cfg_dict, train_cfg)
    trainer.launch_rlg(env)
    trainer.run()
    env.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
