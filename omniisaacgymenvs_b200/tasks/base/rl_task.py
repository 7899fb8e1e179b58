"""RLTask -- RL buffers and the post-physics orchestration (reference: tasks/base/rl_task.py:42-251).

Keeps the reference attribute names (``obs_buf``, ``rew_buf``, ``reset_buf``, ``progress_buf``,
``states_buf``, ``extras``, ``clip_obs``, ``clip_actions``, ``rl_device``, ``control_frequency_inv``,
``randomize_actions``, ``randomize_observations``) because the env wrapper and rl_games read them.
USD cloning / collision filtering / viewport set-up of the reference has no counterpart here:
environments are rows of the state tensors and never interact.
"""
import warnings

import numpy as np
import torch

from omniisaacgymenvs_b200.articulations import grid_env_origins
from omniisaacgymenvs_b200.envs.vec_env_base import Box
from omniisaacgymenvs_b200.utils.domain_randomization.randomize import Randomizer


class BaseTask:
    """The slice of ``omni.isaac.core.tasks.BaseTask`` the RL tasks rely on."""

    def __init__(self, name, offset=None) -> None:
        self._name = name
        self._offset = offset
        self._scene = None

    @property
    def name(self):
        return self._name

    @property
    def device(self):
        return self._device

    def set_up_scene(self, scene) -> None:
        self._scene = scene

    def post_reset(self) -> None:
        pass

    def cleanup(self) -> None:
        pass


class RLTask(BaseTask):
    def __init__(self, name, env, offset=None) -> None:
        super().__init__(name=name, offset=offset)
        self.test = self._cfg.get("test", False)
        self._device = self._cfg["sim_device"]
        self._dr_randomizer = Randomizer(self._sim_config)  # observation / action noise
        print("Task Device:", self._device)

        self.randomize_actions = False
        self.randomize_observations = False

        env_cfg = self._cfg["task"]["env"]
        self.clip_obs = env_cfg.get("clipObservations", np.inf)
        self.clip_actions = env_cfg.get("clipActions", np.inf)
        self.rl_device = self._cfg.get("rl_device", "cuda:0")
        self.control_frequency_inv = env_cfg.get("controlFrequencyInv", 1)
        print("RL device: ", self.rl_device)

        self._env = env
        # env-index sharding across GPUs: this process owns global envs
        # [env_id_offset, env_id_offset + num_envs) of a job with global_num_envs environments.
        self.env_id_offset = int(self._cfg.get("env_id_offset", 0))
        self.global_num_envs = int(self._cfg.get("global_num_envs", self._num_envs + self.env_id_offset))

        if not hasattr(self, "_num_agents"):
            self._num_agents = 1
        if not hasattr(self, "_num_states"):
            self._num_states = 0
        if not hasattr(self, "action_space"):
            self.action_space = Box(np.ones(self.num_actions) * -1.0, np.ones(self.num_actions) * 1.0)
        if not hasattr(self, "observation_space"):
            self.observation_space = Box(np.ones(self.num_observations) * -np.inf, np.ones(self.num_observations) * np.inf)
        if not hasattr(self, "state_space"):
            self.state_space = Box(np.ones(self.num_states) * -np.inf, np.ones(self.num_states) * np.inf)
        self.cleanup()

    def set_up_domain_randomization(self) -> None:
        """Called by ``World.reset`` after ``post_reset``, where the reference's tasks that use the Randomizer call
        it (tasks/shared/in_hand_manipulation.py:206-207).  The reference's locomotion / Anymal / Cartpole tasks never
        do, so a ``domain_randomization`` block in their YAML is inert there; here it is honoured."""
        if self._dr_randomizer.randomize and not getattr(self, "_dr_is_set_up", False):
            self._dr_randomizer.set_up_domain_randomization(self)
            self._dr_is_set_up = True

    # ---- helpers shared by the task classes' C-ABI plumbing -------------------------------------------------
    def _kernel_clip_obs(self) -> float:
        """clip_obs as the observation kernels see it: +inf while observation noise is active, because the
        wrapper's clamp then has to come AFTER the noise (vec_env_rlgames.py:42 vs :70-71) and moves into the noise
        launch.  Survives every rebuild of the parameter structs."""
        return float("inf") if getattr(self, "randomize_observations", False) else float(self.clip_obs)

    def _ensure_reset_counter(self) -> torch.Tensor:
        """Device words {Philox reset counter, CTA ticket} of the CUDA-graph twin (OigeTaskBuffers.reset_counter).
        Allocated ONCE per task: captured graphs and the Randomizer keep the pointer."""
        if getattr(self, "_reset_counter_words", None) is None:
            self._reset_counter_words = torch.zeros(2, dtype=torch.int64, device=self._device)
            self._reset_counter_dev = self._reset_counter_words[:1]  # fill_() on the view leaves the ticket at 0
        return self._reset_counter_dev

    def _invalidate_graphs(self) -> None:
        """Buffers were re-bound: graphs captured with the old pointers must not be replayed.  Goes through the env's
        own disable_* calls, which also hand the Randomizer back its host-side step counter -- the eager step does
        not advance the device word the captured graphs used, so noise drawn from it would repeat every step."""
        env = getattr(self, "_env", None)
        if env is None:
            return
        dropped = False
        if getattr(env, "_graph", None) is not None:
            env.disable_cuda_graph()
            dropped = True
        if getattr(env, "_host_graphs", None):
            env.disable_cuda_graph_host()
            dropped = True
        if dropped and not getattr(self, "_warned_graph_drop", False):
            self._warned_graph_drop = True
            warnings.warn(f"{self.name}: a task buffer was re-assigned while CUDA-graph mode was on; the captured graphs "
                          "were dropped and the env steps eagerly (about 3x the launch overhead) until "
                          "enable_cuda_graph() / enable_cuda_graph_host() is called again", RuntimeWarning, stacklevel=3)

    def cleanup(self) -> None:
        """Allocates the RL buffers (rl_task.py:98-107)."""
        dev = self._device
        self.obs_buf = torch.zeros((self._num_envs, self.num_observations), device=dev, dtype=torch.float)
        self.states_buf = torch.zeros((self._num_envs, self.num_states), device=dev, dtype=torch.float)
        self.rew_buf = torch.zeros(self._num_envs, device=dev, dtype=torch.float)
        self.reset_buf = torch.ones(self._num_envs, device=dev, dtype=torch.long)
        self.progress_buf = torch.zeros(self._num_envs, device=dev, dtype=torch.long)
        self.extras = {}

    def set_up_scene(self, scene) -> None:
        """Computes the env origins (GridCloner layout, rl_task.py:124-126); the ground plane is
        the z = 0 half-space of the contact model."""
        super().set_up_scene(scene)
        if self._sim_config.task_config.get("sim", {}).get("add_ground_plane", True):
            scene.add_default_ground_plane(prim_path="/World/defaultGroundPlane")
        origins = grid_env_origins(self.global_num_envs, self._env_spacing, self.env_id_offset, self._num_envs)
        self._env_pos = torch.tensor(origins, device=self._device, dtype=torch.float)

    @property
    def default_base_env_path(self):
        return "/World/envs"

    @property
    def default_zero_env_path(self):
        return f"{self.default_base_env_path}/env_0"

    @property
    def num_envs(self):
        return self._num_envs

    @property
    def num_actions(self):
        return self._num_actions

    @property
    def num_observations(self):
        return self._num_observations

    @property
    def num_states(self):
        return self._num_states

    @property
    def num_agents(self):
        return self._num_agents

    def get_states(self):
        return self.states_buf

    def get_extras(self):
        return self.extras

    def reset(self):
        """Flags all environments for reset (rl_task.py:218-221).  In place, so that the kernels'
        cached pointer to ``reset_buf`` stays valid."""
        self.reset_buf.fill_(1)

    def pre_physics_step(self, actions):
        pass

    def post_physics_step(self):
        """rl_task.py:231-251."""
        self.progress_buf[:] += 1
        if self._env._world.is_playing():
            self.get_observations()
            self.get_states()
            self.calculate_metrics()
            self.is_done()
            self.get_extras()
        return self.obs_buf, self.rew_buf, self.reset_buf, self.extras
