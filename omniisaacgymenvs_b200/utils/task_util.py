"""Task registry and construction (reference: utils/task_util.py:30-71).  Built here: the locomotion hot path
(Ant, Humanoid), Cartpole and flat-terrain Anymal (N2); the reference's other tasks are out of scope."""
from omniisaacgymenvs_b200.utils.config import SimConfig


def task_map() -> dict:
    from omniisaacgymenvs_b200.tasks.ant import AntLocomotionTask
    from omniisaacgymenvs_b200.tasks.anymal import AnymalTask
    from omniisaacgymenvs_b200.tasks.cartpole import CartpoleTask
    from omniisaacgymenvs_b200.tasks.humanoid import HumanoidLocomotionTask

    return {"Ant": AntLocomotionTask, "Humanoid": HumanoidLocomotionTask, "Cartpole": CartpoleTask,
            "Anymal": AnymalTask}


def initialize_task(config, env, init_sim=True):
    sim_config = SimConfig(config)
    cfg = sim_config.config
    tasks = task_map()
    if cfg["task_name"] not in tasks:
        raise KeyError(f"task '{cfg['task_name']}' is not part of this build; available: {sorted(tasks)}")
    task = tasks[cfg["task_name"]](name=cfg["task_name"], sim_config=sim_config, env=env)
    env.set_task(task=task, sim_params=sim_config.get_physics_params(), backend="torch", init_sim=init_sim)
    return task


def make_env(task: str = "Humanoid", num_envs: int = 4096, device_id: int = 0, seed: int = 42,
             env_id_offset: int = 0, global_num_envs=None, overrides=None):
    """Convenience constructor used by bench.py / tests: config -> VecEnvRLGames + task."""
    from omniisaacgymenvs_b200.envs.vec_env_rlgames import VecEnvRLGames
    from omniisaacgymenvs_b200.utils.config import load_config

    ov = {"num_envs": num_envs, "device_id": device_id, "seed": seed, "rl_device": f"cuda:{device_id}",
          "env_id_offset": env_id_offset,
          "global_num_envs": global_num_envs if global_num_envs is not None else env_id_offset + num_envs}
    ov.update(overrides or {})
    cfg = load_config(task, ov)
    env = VecEnvRLGames(headless=True, sim_device=device_id)
    t = initialize_task(cfg, env)
    return env, t
