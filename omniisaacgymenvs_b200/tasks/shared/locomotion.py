"""LocomotionTask -- shared Ant / Humanoid task logic on the B200 kernels.

Reference: tasks/shared/locomotion.py:45-183 (class) and :190-321 (the four torch.jit functions,
replaced here by the fused K1 kernel).  Method names, buffer names and call order are the
reference's; what changed is where the arithmetic runs:

  post_physics_step   -> ONE launch of K1 (progress += 1, observations, limit cost, reward, done)
  pre_physics_step    -> K3 masked reset on the device (no ``nonzero()`` host sync) + efforts
  reset_idx(env_ids)  -> K3 indexed reset
  fused_step          -> the whole VecEnvRLGames.step as one C-ABI call (K3 -> K2 -> K1)

``get_observations`` / ``calculate_metrics`` / ``is_done`` remain individually callable: the first
runs the fused pass into staging buffers for reward/reset, the other two publish them.
"""
import ctypes as C
import math
from abc import abstractmethod

import numpy as np
import torch

from omniisaacgymenvs_b200 import _lib
from omniisaacgymenvs_b200.tasks.base.rl_task import RLTask
from omniisaacgymenvs_b200.utils.maths import quat_conjugate


class LocomotionTask(RLTask):
    supports_fused_step = True

    def __init__(self, name, env, offset=None) -> None:
        env_cfg = self._task_cfg["env"]
        self._num_envs = env_cfg["numEnvs"]
        self._env_spacing = env_cfg["envSpacing"]
        self._max_episode_length = env_cfg["episodeLength"]
        self.dof_vel_scale = env_cfg["dofVelocityScale"]
        self.angular_velocity_scale = env_cfg["angularVelocityScale"]
        self.contact_force_scale = env_cfg["contactForceScale"]
        self.power_scale = env_cfg["powerScale"]
        self.heading_weight = env_cfg["headingWeight"]
        self.up_weight = env_cfg["upWeight"]
        self.actions_cost_scale = env_cfg["actionsCost"]
        self.energy_cost_scale = env_cfg["energyCost"]
        self.joints_at_limit_cost_scale = env_cfg["jointsAtLimitCost"]
        self.death_cost = env_cfg["deathCost"]
        self.termination_height = env_cfg["terminationHeight"]
        self.alive_reward_scale = env_cfg["alive_reward_scale"]
        self._robots = None
        self._reset_seed = int(self._cfg.get("seed", 42)) & 0xFFFFFFFFFFFFFFFF
        self._reset_counter = 0
        RLTask.__init__(self, name, env)

    # ------------------------------------------------------------------------------------------
    @abstractmethod
    def set_up_scene(self, scene) -> None:
        pass

    @abstractmethod
    def get_robot(self):
        pass

    limit_cost_kind = _lib.LIMIT_COST_HUMANOID

    # ------------------------------------------------------------------------------------------
    # C-ABI plumbing
    # ------------------------------------------------------------------------------------------
    def _build_params(self) -> "_lib.OigeTaskParams":
        p = _lib.OigeTaskParams()
        d = self._robots.num_dof
        p.num_dof, p.num_sensors, p.num_obs = d, self._robots.model.num_sensors, self.num_observations
        p.limit_cost_kind = self.limit_cost_kind
        p.walk_target_uses_z = 1  # upstream compute_rot quirk kept
        p.progress_increment = 1
        p.closed_form_angles = 1  # normalize_angle as a closed-form wrap (1-2 ulp from the literal chain)
        p.dt = self.dt
        p.dof_vel_scale = self.dof_vel_scale
        p.angular_velocity_scale = self.angular_velocity_scale
        p.contact_force_scale = self.contact_force_scale
        p.heading_weight = self.heading_weight
        p.up_weight = self.up_weight
        p.actions_cost_scale = self.actions_cost_scale
        p.energy_cost_scale = self.energy_cost_scale
        p.joints_at_limit_cost_scale = self.joints_at_limit_cost_scale
        p.death_cost = self.death_cost
        p.termination_height = self.termination_height
        p.alive_reward_scale = self.alive_reward_scale
        p.max_episode_length = float(self._max_episode_length)
        p.clip_obs = self._kernel_clip_obs()
        p.power_scale = self.power_scale
        tgt = self.targets[0].tolist()
        inv = self.inv_start_rot[0].tolist()
        b0, b1 = self.basis_vec0[0].tolist(), self.basis_vec1[0].tolist()
        for k in range(3):
            p.targets[k], p.basis_vec0[k], p.basis_vec1[k] = tgt[k], b0[k], b1[k]
        for k in range(4):
            p.inv_start_rot[k] = inv[k]
        lo, hi = self.dof_limits_lower.tolist(), self.dof_limits_upper.tolist()
        mer, gears = self.motor_effort_ratio.tolist(), self.joint_gears.tolist()
        for k in range(d):
            p.dof_limits_lower[k], p.dof_limits_upper[k] = lo[k], hi[k]
            p.motor_effort_ratio[k], p.joint_gears[k] = mer[k], gears[k]
        return p

    def _bind_buffers(self) -> None:
        """(Re)builds the pointer structs; call after any buffer is re-allocated."""
        self._params = self._build_params()
        self._params_noinc = self._build_params()
        self._params_noinc.progress_increment = 0
        self._unit_gear_params = self._build_params()  # World.step feeds efforts, not actions
        for k in range(self._robots.num_dof):
            self._unit_gear_params.joint_gears[k] = 1.0
        self._unit_gear_params.power_scale = 1.0
        self._buffers = _lib.OigeTaskBuffers(
            self.obs_buf.data_ptr(), self.rew_buf.data_ptr(), self.reset_buf.data_ptr(), self.progress_buf.data_ptr(),
            self.potentials.data_ptr(), self.prev_potentials.data_ptr(), self.up_vec.data_ptr(),
            self.heading_vec.data_ptr(), self.actions.data_ptr(), self.episode_stats.data_ptr(), None,
            self.episode_return.data_ptr())
        # CUDA-graph twin: the Philox reset counter lives on the device and is bumped by the step itself
        self._ensure_reset_counter()
        self._buffers_graph = _lib.OigeTaskBuffers(
            self.obs_buf.data_ptr(), self.rew_buf.data_ptr(), self.reset_buf.data_ptr(), self.progress_buf.data_ptr(),
            self.potentials.data_ptr(), self.prev_potentials.data_ptr(), self.up_vec.data_ptr(),
            self.heading_vec.data_ptr(), self.actions.data_ptr(), self.episode_stats.data_ptr(),
            self._reset_counter_dev.data_ptr(), self.episode_return.data_ptr())
        # staging twins for the individually-callable get_observations / calculate_metrics / is_done
        self._staged_rew = torch.zeros_like(self.rew_buf)
        self._staged_reset = torch.zeros_like(self.reset_buf)
        self._staged_progress = torch.zeros_like(self.progress_buf)
        self._staged_buffers = _lib.OigeTaskBuffers(
            self.obs_buf.data_ptr(), self._staged_rew.data_ptr(), self._staged_reset.data_ptr(),
            self._staged_progress.data_ptr(), self.potentials.data_ptr(), self.prev_potentials.data_ptr(),
            self.up_vec.data_ptr(), self.heading_vec.data_ptr(), self.actions.data_ptr(), None, None)
        self._reset_init = _lib.OigeResetInit(
            self.initial_root_pos.data_ptr(), self.initial_root_rot.data_ptr(), self.initial_dof_pos.data_ptr())
        self._bound = (self.obs_buf.data_ptr(), self.rew_buf.data_ptr(), self.reset_buf.data_ptr(),
                       self.progress_buf.data_ptr(), self.actions.data_ptr())

    def _check_bound(self) -> None:
        cur = (self.obs_buf.data_ptr(), self.rew_buf.data_ptr(), self.reset_buf.data_ptr(),
               self.progress_buf.data_ptr(), self.actions.data_ptr())
        if cur != self._bound:  # a caller re-assigned a buffer attribute (the reference does in reset())
            self._bind_buffers()
            self._invalidate_graphs()

    def _stream(self) -> int:
        return _lib.current_stream_ptr(self._device)

    # ------------------------------------------------------------------------------------------
    # reference surface
    # ------------------------------------------------------------------------------------------
    def get_observations(self) -> dict:
        """locomotion.py:80-101.  Runs the fused pass; reward / reset land in staging buffers."""
        self._check_bound()
        self._staged_reset.copy_(self.reset_buf)
        self._staged_progress.copy_(self.progress_buf)
        lib = _lib.load()
        rc = lib.oige_obs_reward_reset_f32(C.byref(self._params_noinc), C.byref(self._robots.state_struct),
                                           C.byref(self._staged_buffers), self._num_envs, self._stream())
        _lib.check(rc, "oige_obs_reward_reset_f32")
        return {self._robots.name: {"obs_buf": self.obs_buf}}

    def calculate_metrics(self) -> None:
        """locomotion.py:173-178 -- publishes the reward computed by the fused pass."""
        self.rew_buf.copy_(self._staged_rew)

    def is_done(self) -> None:
        """locomotion.py:180-183 -- publishes the reset flags computed by the fused pass."""
        self.reset_buf.copy_(self._staged_reset)

    def _hooks_overridden(self) -> bool:
        cls = type(self)
        return (cls.get_observations is not LocomotionTask.get_observations
                or cls.calculate_metrics is not LocomotionTask.calculate_metrics
                or cls.is_done is not LocomotionTask.is_done)

    def post_physics_step(self):
        """rl_task.py:231-251 as one K1 launch (falls back to the hook-by-hook sequence if a
        subclass overrides one of the hooks)."""
        if self._hooks_overridden():
            return RLTask.post_physics_step(self)
        self._check_bound()
        if self._env._world.is_playing():
            lib = _lib.load()
            rc = lib.oige_obs_reward_reset_f32(C.byref(self._params), C.byref(self._robots.state_struct),
                                               C.byref(self._buffers), self._num_envs, self._stream())
            _lib.check(rc, "oige_obs_reward_reset_f32")
        else:
            self.progress_buf[:] += 1
        return self.obs_buf, self.rew_buf, self.reset_buf, self.extras

    def pre_physics_step(self, actions) -> None:
        """locomotion.py:103-114 without the host sync: the reset mask is consumed on the device."""
        self._check_bound()
        lib = _lib.load()
        self._reset_counter += 1
        rc = lib.oige_reset_masked_f32(C.byref(self._params), C.byref(self._robots.state_struct),
                                       C.byref(self._buffers), C.byref(self._reset_init), None, None,
                                       self._reset_seed, self._reset_counter, self.env_id_offset, self._num_envs,
                                       self._stream())
        _lib.check(rc, "oige_reset_masked_f32")
        self.actions.copy_(actions.to(self._device))
        forces = self.actions * self.joint_gears * self.power_scale
        indices = torch.arange(self._robots.count, dtype=torch.int32, device=self._device)
        self._robots.set_joint_efforts(forces, indices=indices)

    def reset_idx(self, env_ids, rand_pos=None, rand_vel=None):
        """locomotion.py:116-145.  ``rand_pos`` / ``rand_vel`` ([k,D] raw U[0,1) draws) reproduce
        the reference's torch.rand stream when given; otherwise the kernel draws Philox numbers
        keyed by (seed, global env id, reset counter)."""
        self._check_bound()
        env_ids = env_ids.to(device=self._device, dtype=torch.int64).contiguous()
        k = int(env_ids.numel())
        if k == 0:
            return
        lib = _lib.load()
        self._reset_counter += 1
        rp = _lib.dptr(rand_pos, torch.float32, "rand_pos")
        rv = _lib.dptr(rand_vel, torch.float32, "rand_vel")
        rc = lib.oige_reset_idx_f32(C.byref(self._params), C.byref(self._robots.state_struct), C.byref(self._buffers),
                                    C.byref(self._reset_init), env_ids.data_ptr(), k, rp, rv, self._reset_seed,
                                    self._reset_counter, self.env_id_offset, self._stream())
        _lib.check(rc, "oige_reset_idx_f32")
        self.episode_return[env_ids] = 0.0  # a hand-reset env starts a new episode: its partial return is dropped

    def reset(self):
        """rl_task.py:218-221, plus: the partial returns of the episodes cut short here are not published."""
        super().reset()
        if hasattr(self, "episode_return"):
            self.episode_return.zero_()

    def post_reset(self):
        """locomotion.py:147-171."""
        self._robots = self.get_robot()
        self.initial_root_pos, self.initial_root_rot = self._robots.get_world_poses()
        self.initial_dof_pos = self._robots.get_joint_positions()

        dev, n = self._device, self.num_envs
        self.start_rotation = torch.tensor([1, 0, 0, 0], device=dev, dtype=torch.float32)
        self.up_vec = torch.tensor([0, 0, 1], dtype=torch.float32, device=dev).repeat((n, 1))
        self.heading_vec = torch.tensor([1, 0, 0], dtype=torch.float32, device=dev).repeat((n, 1))
        self.inv_start_rot = quat_conjugate(self.start_rotation).repeat((n, 1))
        self.basis_vec0 = self.heading_vec.clone()
        self.basis_vec1 = self.up_vec.clone()
        self.targets = torch.tensor([1000, 0, 0], dtype=torch.float32, device=dev).repeat((n, 1))
        self.target_dirs = torch.tensor([1, 0, 0], dtype=torch.float32, device=dev).repeat((n, 1))
        self.dt = 1.0 / 60.0
        self.potentials = torch.tensor([-1000.0 / self.dt], dtype=torch.float32, device=dev).repeat(n)
        self.prev_potentials = self.potentials.clone()
        self.actions = torch.zeros((n, self.num_actions), device=dev)
        # episode statistics accumulated by K1 / K2-F (sum of the returns of finished episodes, #done, sum len at done,
        # #env-steps), striped; episode_return is each env's running return, published when its episode ends
        self.episode_stats = torch.zeros((_lib.STATS_SLOTS, 4), device=dev, dtype=torch.float64)
        self.episode_return = torch.zeros(n, device=dev, dtype=torch.float32)
        # landing buffer of poll_episode_extras, pinned HERE (cudaHostAlloc synchronises the device and takes
        # milliseconds: not something to do lazily inside a step)
        self._ep_host = torch.zeros(4, dtype=torch.float64).pin_memory()
        self._ep_prev = torch.zeros(4, dtype=torch.float64)
        self._ep_event = None
        self._ep_step = 0
        # first use of the reduction / copy / event paths loads their CUDA modules (milliseconds): do it now
        self._ep_host.copy_(self.episode_stats.sum(dim=0), non_blocking=True)
        warm = torch.cuda.Event()
        warm.record(torch.cuda.current_stream(dev))
        warm.synchronize()
        self._bind_buffers()

        indices = torch.arange(self._robots.count, dtype=torch.int64, device=dev)
        self.reset_idx(indices)

    # ------------------------------------------------------------------------------------------
    # whole-step entry points
    # ------------------------------------------------------------------------------------------
    def fused_step_enabled(self) -> bool:
        cls = type(self)
        return (not self._hooks_overridden()
                and cls.pre_physics_step is LocomotionTask.pre_physics_step
                and cls.post_physics_step is LocomotionTask.post_physics_step
                and cls.reset_idx is LocomotionTask.reset_idx
                and self._env._world.is_playing())

    def fused_step(self, actions: torch.Tensor, physics_dt: float, graph_mode: bool = False,
                   clip_actions=None) -> None:
        """VecEnvRLGames.step (envs/vec_env_rlgames.py:56-78) as one C-ABI call.  ``graph_mode`` takes the
        reset counter from device memory so the call can be captured in a CUDA graph."""
        self._check_bound()
        a = actions
        if a.dtype != torch.float32 or not a.is_contiguous():
            a = a.float().contiguous()
        self._reset_counter += 1
        lib = _lib.load()
        rc = lib.oige_env_step_f32(
            self._robots.model_handle, C.byref(self._params), C.byref(self._robots.state_struct),
            C.byref(self._buffers_graph if graph_mode else self._buffers), C.byref(self._reset_init),
            _lib.dptr(a, torch.float32, "actions"),
            float(self.clip_actions if clip_actions is None else clip_actions), self._reset_seed, self._reset_counter, self.env_id_offset, self._num_envs,
            int(self.control_frequency_inv), float(physics_dt), self._stream())
        _lib.check(rc, "oige_env_step_f32")

    def fused_step_host(self, actions_host, obs_host, rew_host, reset_host, physics_dt: float,
                        graph_mode: bool = False, run: bool = True) -> None:
        self._check_bound()
        if not hasattr(self, "_actions_staging"):
            self._actions_staging = torch.empty_like(self.actions)
        for t, nm in ((actions_host, "actions_host"), (obs_host, "obs_host"), (rew_host, "rew_host"),
                      (reset_host, "reset_host")):
            if t.is_cuda or not t.is_contiguous():
                raise RuntimeError(f"{nm} must be a contiguous host tensor")
        n = self._num_envs if run else 0  # n = 0: only creates the library's copy/compute streams
        self._reset_counter += 1 if run else 0
        lib = _lib.load()
        rc = lib.oige_env_step_host_f32(
            self._robots.model_handle, C.byref(self._params), C.byref(self._robots.state_struct),
            C.byref(self._buffers_graph if graph_mode else self._buffers), C.byref(self._reset_init), actions_host.data_ptr(),
            self._actions_staging.data_ptr(), obs_host.data_ptr(), rew_host.data_ptr(), reset_host.data_ptr(),
            float(self.clip_actions), self._reset_seed, self._reset_counter, self.env_id_offset, n,
            int(self.control_frequency_inv), float(physics_dt), self._stream())
        _lib.check(rc, "oige_env_step_host_f32")

    def poll_episode_extras(self, every: int = 16) -> None:
        """Puts the episode statistics where the reference's observer looks (``infos["episode"]``,
        utils/rlgames/rlgames_utils.py:51-62,67-91) WITHOUT a per-step sync: every ``every`` steps the striped
        accumulators are summed on the device and copied to pinned host memory asynchronously; the copy enqueued
        earlier is published as the window's mean return (sum of the returns of the episodes that FINISHED in the window,
        each accumulated per env in ``episode_return`` by the kernel, over their count) / mean length / episode count
        once its event has completed
        (polled with ``query()``, never waited for: a host that runs ahead of the device just tries again later).  ``extras["episode"]`` is present only in the step that publishes a window (the observer
        appends whatever it finds, so a stale entry would be counted again)."""
        self.extras.pop("episode", None)
        self._ep_step += 1
        if self._ep_step % max(int(every), 1):
            return
        dev = torch.device(self._device)
        if self._ep_event is not None:
            if not self._ep_event.query():  # the host runs ahead of the device: never wait, try again next time
                return
            tot = self._ep_host.clone()
            win = tot - self._ep_prev
            if bool((win < 0).any()):  # the accumulators were read-and-reset in between (read_episode_stats)
                win = tot
            self._ep_prev = tot
            if float(win[1]) > 0:
                self.extras["episode"] = {"mean_return": torch.tensor(float(win[0] / win[1])),
                                          "mean_length": torch.tensor(float(win[2] / win[1])),
                                          "count": torch.tensor(float(win[1]))}
        self._ep_host.copy_(self.episode_stats.sum(dim=0), non_blocking=True)
        self._ep_event = torch.cuda.Event()
        self._ep_event.record(torch.cuda.current_stream(dev))

    def read_episode_stats(self, reset: bool = True) -> torch.Tensor:
        """[4] float64 on the device: sum(reward), #episodes finished, sum(length at finish), #env-steps."""
        s = self.episode_stats.sum(dim=0)
        if reset:
            self.episode_stats.zero_()
        return s
