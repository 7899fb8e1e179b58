"""CartpoleTask (reference: tasks/cartpole.py:40-162) on the B200 kernels: 4 observations, 1 action, a
3-link welded-base tree (rail, cart on a prismatic joint, pole on a hinge) stepped by the SAME articulation
kernel as the locomotion robots.  Method names, buffer names and call order are the reference's:

  post_physics_step   -> one launch: progress += 1, observations, reward, done (cartpole.py:80-99,143-162)
  pre_physics_step    -> one launch: device-side masked reset (no ``nonzero()`` host sync) + efforts (:101-108)
  reset_idx(env_ids)  -> indexed reset kernel (:110-130)
  fused_step          -> the whole VecEnvRLGames.step as one C-ABI call (three launches)

``obs_buf`` holds the observations already clamped to +-clipObservations (what ``VecEnvRLGames.step``
returns, vec_env_rlgames.py:42); reward and done are evaluated on the unclamped values like the reference.
"""
import ctypes as C

import torch

from omniisaacgymenvs_b200 import _lib
from omniisaacgymenvs_b200.articulations import ArticulationView
from omniisaacgymenvs_b200.robots.model import build_model
from omniisaacgymenvs_b200.tasks.base.rl_task import RLTask
from omniisaacgymenvs_b200.tasks.humanoid import _solver_overrides


class CartpoleTask(RLTask):
    supports_fused_step = True

    def __init__(self, name, sim_config, env, offset=None) -> None:
        self._sim_config = sim_config
        self._cfg = sim_config.config
        self._task_cfg = sim_config.task_config

        self._num_envs = self._task_cfg["env"]["numEnvs"]
        self._env_spacing = self._task_cfg["env"]["envSpacing"]
        self._cartpole_positions = torch.tensor([0.0, 0.0, 2.0])

        self._reset_dist = self._task_cfg["env"]["resetDist"]
        self._max_push_effort = self._task_cfg["env"]["maxEffort"]
        self._max_episode_length = 500

        self._num_observations = 4
        self._num_actions = 1
        self._robots = None
        self._reset_seed = int(self._cfg.get("seed", 42)) & 0xFFFFFFFFFFFFFFFF
        self._reset_counter = 0
        RLTask.__init__(self, name, env)

    # ---- scene ------------------------------------------------------------------------------------------
    def set_up_scene(self, scene) -> None:
        self.get_cartpole()
        RLTask.set_up_scene(self, scene)
        self._cartpoles = ArticulationView(self._model, self._num_envs, self._device, self._env_pos,
                                           name="cartpole_view")
        scene.add(self._cartpoles)

    def get_cartpole(self):
        self._model = build_model("Cartpole", spawn_height=float(self._cartpole_positions[2]),
                                  **_solver_overrides(self._sim_config, "Cartpole"))

    def get_robot(self):
        return self._cartpoles

    # ---- C-ABI plumbing -----------------------------------------------------------------------------------
    def _build_params(self, progress_increment: int) -> "_lib.OigeCartpoleParams":
        p = _lib.OigeCartpoleParams()
        p.cart_dof, p.pole_dof = int(self._cart_dof_idx), int(self._pole_dof_idx)
        p.progress_increment = progress_increment
        p.reset_dist = float(self._reset_dist)
        p.max_push_effort = float(self._max_push_effort)
        p.max_episode_length = float(self._max_episode_length)
        p.clip_obs = self._kernel_clip_obs()
        return p

    def _bind_buffers(self) -> None:
        v = self._cartpoles
        self._params = self._build_params(1)
        self._params_noinc = self._build_params(0)
        self._unit_gear_params = _lib.OigeTaskParams()  # World.step feeds the efforts written by pre_physics_step
        self._unit_gear_params.num_dof = v.num_dof
        for k in range(v.num_dof):
            self._unit_gear_params.joint_gears[k] = 1.0
        self._unit_gear_params.power_scale = 1.0
        ptrs = (self.obs_buf.data_ptr(), self.rew_buf.data_ptr(), self.reset_buf.data_ptr(),
                self.progress_buf.data_ptr(), v.dof_efforts.data_ptr())
        self._buffers = _lib.OigeCartpoleBuffers(*ptrs, self.episode_stats.data_ptr(), None)
        self._ensure_reset_counter()
        self._buffers_graph = _lib.OigeCartpoleBuffers(*ptrs, self.episode_stats.data_ptr(),
                                                       self._reset_counter_dev.data_ptr())
        # staging twins for the individually-callable get_observations / calculate_metrics / is_done
        self._staged_rew = torch.zeros_like(self.rew_buf)
        self._staged_reset = torch.zeros_like(self.reset_buf)
        self._staged_progress = torch.zeros_like(self.progress_buf)
        self._staged_buffers = _lib.OigeCartpoleBuffers(
            self.obs_buf.data_ptr(), self._staged_rew.data_ptr(), self._staged_reset.data_ptr(),
            self._staged_progress.data_ptr(), v.dof_efforts.data_ptr(), None, None)
        self._bound = ptrs

    def _check_bound(self) -> None:
        cur = (self.obs_buf.data_ptr(), self.rew_buf.data_ptr(), self.reset_buf.data_ptr(),
               self.progress_buf.data_ptr(), self._cartpoles.dof_efforts.data_ptr())
        if cur != self._bound:
            self._bind_buffers()
            self._invalidate_graphs()

    def _stream(self) -> int:
        return _lib.current_stream_ptr(self._device)

    # ---- reference surface --------------------------------------------------------------------------------
    def get_observations(self) -> dict:
        """cartpole.py:80-99.  Runs the fused pass; reward / reset land in staging buffers."""
        self._check_bound()
        self._staged_progress.copy_(self.progress_buf)
        rc = _lib.load().oige_cartpole_obs_reward_reset_f32(
            C.byref(self._params_noinc), C.byref(self._cartpoles.state_struct), C.byref(self._staged_buffers),
            self._num_envs, self._stream())
        _lib.check(rc, "oige_cartpole_obs_reward_reset_f32")
        return {self._cartpoles.name: {"obs_buf": self.obs_buf}}

    def calculate_metrics(self) -> None:
        """cartpole.py:143-153 -- publishes the reward computed by the fused pass."""
        self.rew_buf.copy_(self._staged_rew)

    def is_done(self) -> None:
        """cartpole.py:155-162 -- publishes the reset flags computed by the fused pass."""
        self.reset_buf.copy_(self._staged_reset)

    def _hooks_overridden(self) -> bool:
        cls = type(self)
        return (cls.get_observations is not CartpoleTask.get_observations
                or cls.calculate_metrics is not CartpoleTask.calculate_metrics
                or cls.is_done is not CartpoleTask.is_done)

    def post_physics_step(self):
        """rl_task.py:231-251 as one launch (hook-by-hook if a subclass overrides a hook)."""
        if self._hooks_overridden():
            return RLTask.post_physics_step(self)
        self._check_bound()
        if self._env._world.is_playing():
            rc = _lib.load().oige_cartpole_obs_reward_reset_f32(
                C.byref(self._params), C.byref(self._cartpoles.state_struct), C.byref(self._buffers), self._num_envs,
                self._stream())
            _lib.check(rc, "oige_cartpole_obs_reward_reset_f32")
        else:
            self.progress_buf[:] += 1
        return self.obs_buf, self.rew_buf, self.reset_buf, self.extras

    def pre_physics_step(self, actions) -> None:
        """cartpole.py:101-108 without the host sync: the reset mask is consumed on the device and the efforts
        are written straight into the view's effort buffer (what set_joint_efforts would have received)."""
        self._check_bound()
        a = actions.to(self._device)
        if a.dtype != torch.float32 or not a.is_contiguous():
            a = a.float().contiguous()
        self._reset_counter += 1
        rc = _lib.load().oige_cartpole_reset_masked_f32(
            C.byref(self._params), C.byref(self._cartpoles.state_struct), C.byref(self._buffers),
            _lib.dptr(a, torch.float32, "actions"), float("inf"), None, self._reset_seed, self._reset_counter,
            self.env_id_offset, self._num_envs, self._stream())
        _lib.check(rc, "oige_cartpole_reset_masked_f32")

    def reset_idx(self, env_ids, rand=None):
        """cartpole.py:110-130.  ``rand`` ([k,4] raw U[0,1) draws in the reference's draw order) reproduces the
        reference's torch.rand stream when given; otherwise Philox keyed by (seed, global env id, counter)."""
        self._check_bound()
        env_ids = env_ids.to(device=self._device, dtype=torch.int64).contiguous()
        k = int(env_ids.numel())
        if k == 0:
            return
        self._reset_counter += 1
        rc = _lib.load().oige_cartpole_reset_idx_f32(
            C.byref(self._params), C.byref(self._cartpoles.state_struct), C.byref(self._buffers), env_ids.data_ptr(), k,
            _lib.dptr(rand, torch.float32, "rand"), self._reset_seed, self._reset_counter, self.env_id_offset,
            self._stream())
        _lib.check(rc, "oige_cartpole_reset_idx_f32")

    def post_reset(self):
        """cartpole.py:132-137."""
        self._robots = self._cartpoles
        self._cart_dof_idx = self._cartpoles.get_dof_index("cartJoint")
        self._pole_dof_idx = self._cartpoles.get_dof_index("poleJoint")
        self.episode_stats = torch.zeros((_lib.STATS_SLOTS, 4), device=self._device, dtype=torch.float64)
        self._bind_buffers()
        indices = torch.arange(self._cartpoles.count, dtype=torch.int64, device=self._device)
        self.reset_idx(indices)

    # ---- whole-step entry points --------------------------------------------------------------------------
    def fused_step_enabled(self) -> bool:
        cls = type(self)
        return (not self._hooks_overridden()
                and cls.pre_physics_step is CartpoleTask.pre_physics_step
                and cls.post_physics_step is CartpoleTask.post_physics_step
                and cls.reset_idx is CartpoleTask.reset_idx
                and self._env._world.is_playing())

    def fused_step(self, actions: torch.Tensor, physics_dt: float, graph_mode: bool = False,
                   clip_actions=None) -> None:
        """VecEnvRLGames.step (envs/vec_env_rlgames.py:56-78) as one C-ABI call; ``graph_mode`` takes the reset
        counter from device memory so the call can be captured in a CUDA graph."""
        self._check_bound()
        a = actions
        if a.dtype != torch.float32 or not a.is_contiguous():
            a = a.float().contiguous()
        self._reset_counter += 1
        rc = _lib.load().oige_cartpole_env_step_f32(
            self._cartpoles.model_handle, C.byref(self._params), C.byref(self._cartpoles.state_struct),
            C.byref(self._buffers_graph if graph_mode else self._buffers), _lib.dptr(a, torch.float32, "actions"),
            float(self.clip_actions if clip_actions is None else clip_actions), self._reset_seed, self._reset_counter, self.env_id_offset, self._num_envs,
            int(self.control_frequency_inv), float(physics_dt), self._stream())
        _lib.check(rc, "oige_cartpole_env_step_f32")

    def read_episode_stats(self, reset: bool = True) -> torch.Tensor:
        """[4] float64 on the device: sum(reward), #episodes finished, sum(length at finish), #env-steps."""
        s = self.episode_stats.sum(dim=0)
        if reset:
            self.episode_stats.zero_()
        return s
