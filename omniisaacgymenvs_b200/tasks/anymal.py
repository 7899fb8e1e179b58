"""AnymalTask, flat terrain (reference: tasks/anymal.py:43-293) on the B200 kernels: 48 observations, 12 actions,
PD position-target actuation (``set_joint_position_targets``) through the articulation kernel's implicit drive.
Method names, buffer names and call order are the reference's:

  post_physics_step   -> one launch: progress += 1, observations, reward, fall / time-out (anymal.py:137-179,252-293)
  pre_physics_step    -> one launch: device-side masked reset (no ``nonzero()`` host sync), action store, target
                         integration and clamp (anymal.py:181-190)
  reset_idx(env_ids)  -> indexed reset kernel (anymal.py:192-227)
  fused_step          -> the whole VecEnvRLGames.step as one C-ABI call (three launches)

``obs_buf`` holds the observations already clamped to +-clipObservations (what ``VecEnvRLGames.step`` returns).
"""
import ctypes as C

import torch

from omniisaacgymenvs_b200 import _lib
from omniisaacgymenvs_b200.articulations import ArticulationView
from omniisaacgymenvs_b200.robots.model import build_model
from omniisaacgymenvs_b200.tasks.base.rl_task import RLTask
from omniisaacgymenvs_b200.tasks.humanoid import _solver_overrides


class AnymalView(ArticulationView):
    """robots/articulations/views/anymal_view.py:36-66 (the knee / base RigidPrimViews have no counterpart: the base
    pose is the articulation root pose)."""

    def is_base_below_threshold(self, threshold, ground_heights):
        base_pos, _ = self.get_world_poses()
        base_heights = base_pos[:, 2]
        base_heights -= ground_heights
        return base_heights[:] < threshold


class AnymalTask(RLTask):
    supports_fused_step = True

    def __init__(self, name, sim_config, env, offset=None) -> None:
        self._sim_config = sim_config
        self._cfg = sim_config.config
        self._task_cfg = sim_config.task_config
        env_cfg = self._task_cfg["env"]

        # normalization
        self.lin_vel_scale = env_cfg["learn"]["linearVelocityScale"]
        self.ang_vel_scale = env_cfg["learn"]["angularVelocityScale"]
        self.dof_pos_scale = env_cfg["learn"]["dofPositionScale"]
        self.dof_vel_scale = env_cfg["learn"]["dofVelocityScale"]
        self.action_scale = env_cfg["control"]["actionScale"]

        # reward scales
        self.rew_scales = {
            "lin_vel_xy": env_cfg["learn"]["linearVelocityXYRewardScale"],
            "ang_vel_z": env_cfg["learn"]["angularVelocityZRewardScale"],
            "lin_vel_z": env_cfg["learn"]["linearVelocityZRewardScale"],
            "joint_acc": env_cfg["learn"]["jointAccRewardScale"],
            "action_rate": env_cfg["learn"]["actionRateRewardScale"],
            "cosmetic": env_cfg["learn"]["cosmeticRewardScale"],
        }

        # command ranges
        self.command_x_range = env_cfg["randomCommandVelocityRanges"]["linear_x"]
        self.command_y_range = env_cfg["randomCommandVelocityRanges"]["linear_y"]
        self.command_yaw_range = env_cfg["randomCommandVelocityRanges"]["yaw"]

        # base init state
        bis = env_cfg["baseInitState"]
        self.base_init_state = list(bis["pos"]) + list(bis["rot"]) + list(bis["vLinear"]) + list(bis["vAngular"])
        self.named_default_joint_angles = env_cfg["defaultJointAngles"]

        # other
        self.dt = 1 / 60
        self.max_episode_length_s = env_cfg["learn"]["episodeLength_s"]
        self.max_episode_length = int(self.max_episode_length_s / self.dt + 0.5)
        self.Kp = env_cfg["control"]["stiffness"]
        self.Kd = env_cfg["control"]["damping"]
        for key in self.rew_scales.keys():
            self.rew_scales[key] *= self.dt

        self._num_envs = env_cfg["numEnvs"]
        self._anymal_translation = torch.tensor([0.0, 0.0, 0.62])
        self._env_spacing = env_cfg["envSpacing"]
        self._num_observations = 48
        self._num_actions = 12
        self._robots = None
        self._reset_seed = int(self._cfg.get("seed", 42)) & 0xFFFFFFFFFFFFFFFF
        self._reset_counter = 0
        RLTask.__init__(self, name, env)

    # ---- scene ------------------------------------------------------------------------------------------
    def set_up_scene(self, scene) -> None:
        self.get_anymal()
        RLTask.set_up_scene(self, scene)
        self._anymals = AnymalView(self._model, self._num_envs, self._device, self._env_pos, name="anymalview")
        scene.add(self._anymals)

    def get_anymal(self):
        """anymal.py:116-135: the model table carries the drive set by ``set_drive`` (400 / 40 per degree, max force
        1000); the default joint angles come from the YAML in DoF order."""
        self._model = build_model("Anymal", spawn_height=float(self._anymal_translation[2]),
                                  **_solver_overrides(self._sim_config, "Anymal"))
        self.default_dof_pos = torch.zeros((self.num_envs, 12), dtype=torch.float, device=self.device)
        for i, name in enumerate(self._model.dof_names):
            self.default_dof_pos[:, i] = self.named_default_joint_angles[name]

    def get_robot(self):
        return self._anymals

    # ---- C-ABI plumbing -----------------------------------------------------------------------------------
    def _build_params(self, progress_increment: int) -> "_lib.OigeAnymalParams":
        p = _lib.OigeAnymalParams()
        p.progress_increment = progress_increment
        p.lin_vel_scale, p.ang_vel_scale = self.lin_vel_scale, self.ang_vel_scale
        p.dof_pos_scale, p.dof_vel_scale = self.dof_pos_scale, self.dof_vel_scale
        p.action_scale, p.dt = self.action_scale, self.dt
        p.rew_lin_vel_xy, p.rew_ang_vel_z = self.rew_scales["lin_vel_xy"], self.rew_scales["ang_vel_z"]
        p.rew_lin_vel_z, p.rew_joint_acc = self.rew_scales["lin_vel_z"], self.rew_scales["joint_acc"]
        p.rew_action_rate, p.rew_cosmetic = self.rew_scales["action_rate"], self.rew_scales["cosmetic"]
        for k in range(2):
            p.command_x[k], p.command_y[k] = float(self.command_x_range[k]), float(self.command_y_range[k])
            p.command_yaw[k] = float(self.command_yaw_range[k])
        p.fallen_threshold = 0.51  # anymal.py:284
        p.max_episode_length = float(self.max_episode_length)
        p.clip_obs = self._kernel_clip_obs()
        dflt = self.default_dof_pos[0].tolist()
        lo, hi = self.anymal_dof_lower_limits.tolist(), self.anymal_dof_upper_limits.tolist()
        for k in range(12):
            p.default_dof_pos[k], p.dof_limits_lower[k], p.dof_limits_upper[k] = dflt[k], lo[k], hi[k]
        return p

    def _buffer_ptrs(self):
        return (self.obs_buf.data_ptr(), self.rew_buf.data_ptr(), self.reset_buf.data_ptr(),
                self.progress_buf.data_ptr(), self.actions.data_ptr(), self.last_actions.data_ptr(),
                self.last_dof_vel.data_ptr(), self.commands.data_ptr(), self.current_targets.data_ptr())

    def _bind_buffers(self) -> None:
        self._params = self._build_params(1)
        self._params_noinc = self._build_params(0)
        self._unit_gear_params = _lib.OigeTaskParams()  # World.step: no effort actuation, PD targets only
        self._unit_gear_params.num_dof = 12
        self._unit_gear_params.power_scale = 1.0
        ptrs = self._buffer_ptrs()
        self._buffers = _lib.OigeAnymalBuffers(*ptrs, self.episode_stats.data_ptr(), None)
        self._ensure_reset_counter()
        self._buffers_graph = _lib.OigeAnymalBuffers(*ptrs, self.episode_stats.data_ptr(),
                                                     self._reset_counter_dev.data_ptr())
        # staging twins for the individually-callable get_observations / calculate_metrics / is_done
        self._staged_rew = torch.zeros_like(self.rew_buf)
        self._staged_reset = torch.zeros_like(self.reset_buf)
        self._staged_progress = torch.zeros_like(self.progress_buf)
        self._staged_last_actions = torch.zeros_like(self.last_actions)
        self._staged_last_dof_vel = torch.zeros_like(self.last_dof_vel)
        self._staged_buffers = _lib.OigeAnymalBuffers(
            self.obs_buf.data_ptr(), self._staged_rew.data_ptr(), self._staged_reset.data_ptr(),
            self._staged_progress.data_ptr(), self.actions.data_ptr(), self._staged_last_actions.data_ptr(),
            self._staged_last_dof_vel.data_ptr(), self.commands.data_ptr(), self.current_targets.data_ptr(), None, None)
        self._reset_init = _lib.OigeResetInit(self.initial_root_pos.data_ptr(), self.initial_root_rot.data_ptr(), None)
        self._bound = ptrs

    def _check_bound(self) -> None:
        if self._buffer_ptrs() != self._bound:
            self._bind_buffers()
            self._invalidate_graphs()

    def _stream(self) -> int:
        return _lib.current_stream_ptr(self._device)

    # ---- reference surface --------------------------------------------------------------------------------
    def get_observations(self) -> dict:
        """anymal.py:137-179.  Runs the fused pass; reward / reset / last_* land in staging buffers."""
        self._check_bound()
        self._staged_progress.copy_(self.progress_buf)
        self._staged_last_actions.copy_(self.last_actions)
        self._staged_last_dof_vel.copy_(self.last_dof_vel)
        rc = _lib.load().oige_anymal_obs_reward_reset_f32(
            C.byref(self._params_noinc), C.byref(self._anymals.state_struct), C.byref(self._staged_buffers),
            self._num_envs, self._stream())
        _lib.check(rc, "oige_anymal_obs_reward_reset_f32")
        return {self._anymals.name: {"obs_buf": self.obs_buf}}

    def calculate_metrics(self) -> None:
        """anymal.py:252-287 -- publishes reward, last_actions, last_dof_vel and fallen_over of the fused pass."""
        self.rew_buf.copy_(self._staged_rew)
        self.last_actions.copy_(self._staged_last_actions)
        self.last_dof_vel.copy_(self._staged_last_dof_vel)
        self.fallen_over = self._anymals.is_base_below_threshold(threshold=0.51, ground_heights=0.0)

    def is_done(self) -> None:
        """anymal.py:289-293 -- publishes the reset flags computed by the fused pass."""
        self.reset_buf.copy_(self._staged_reset)

    def _hooks_overridden(self) -> bool:
        cls = type(self)
        return (cls.get_observations is not AnymalTask.get_observations
                or cls.calculate_metrics is not AnymalTask.calculate_metrics
                or cls.is_done is not AnymalTask.is_done)

    def post_physics_step(self):
        """rl_task.py:231-251 as one launch (hook-by-hook if a subclass overrides a hook)."""
        if self._hooks_overridden():
            return RLTask.post_physics_step(self)
        self._check_bound()
        if self._env._world.is_playing():
            rc = _lib.load().oige_anymal_obs_reward_reset_f32(
                C.byref(self._params), C.byref(self._anymals.state_struct), C.byref(self._buffers), self._num_envs,
                self._stream())
            _lib.check(rc, "oige_anymal_obs_reward_reset_f32")
        else:
            self.progress_buf[:] += 1
        return self.obs_buf, self.rew_buf, self.reset_buf, self.extras

    def pre_physics_step(self, actions) -> None:
        """anymal.py:181-190 without the host sync; the integrated targets are written straight into the view's
        position-target buffer (what set_joint_position_targets would have received)."""
        self._check_bound()
        a = actions.to(self._device)
        if a.dtype != torch.float32 or not a.is_contiguous():
            a = a.float().contiguous()
        self._reset_counter += 1
        rc = _lib.load().oige_anymal_reset_masked_f32(
            C.byref(self._params), C.byref(self._anymals.state_struct), C.byref(self._buffers), C.byref(self._reset_init),
            _lib.dptr(a, torch.float32, "actions"), float("inf"), None, None, self._reset_seed, self._reset_counter,
            self.env_id_offset, self._num_envs, self._stream())
        _lib.check(rc, "oige_anymal_reset_masked_f32")

    def reset_idx(self, env_ids, rand_vel=None, rand_cmd=None):
        """anymal.py:192-227.  ``rand_vel`` [k,12] / ``rand_cmd`` [k,3] (raw U[0,1) draws in the reference's draw
        order) reproduce the reference's torch.rand stream; otherwise Philox keyed by (seed, global env id, counter)."""
        self._check_bound()
        env_ids = env_ids.to(device=self._device, dtype=torch.int64).contiguous()
        k = int(env_ids.numel())
        if k == 0:
            return
        self._reset_counter += 1
        rc = _lib.load().oige_anymal_reset_idx_f32(
            C.byref(self._params), C.byref(self._anymals.state_struct), C.byref(self._buffers), C.byref(self._reset_init),
            env_ids.data_ptr(), k, _lib.dptr(rand_vel, torch.float32, "rand_vel"),
            _lib.dptr(rand_cmd, torch.float32, "rand_cmd"), self._reset_seed, self._reset_counter, self.env_id_offset,
            self._stream())
        _lib.check(rc, "oige_anymal_reset_idx_f32")

    def post_reset(self):
        """anymal.py:229-250."""
        self._robots = self._anymals
        dev, n = self._device, self._num_envs
        self.initial_root_pos, self.initial_root_rot = self._anymals.get_world_poses()
        # the view's position-target buffer IS current_targets: the pre-physics kernel integrates in place
        self._anymals.set_joint_position_targets(self.default_dof_pos)
        self.current_targets = self._anymals.dof_pos_targets

        dof_limits = self._anymals.get_dof_limits()
        self.anymal_dof_lower_limits = dof_limits[0, :, 0].to(device=dev)
        self.anymal_dof_upper_limits = dof_limits[0, :, 1].to(device=dev)

        self.commands = torch.zeros(n, 3, dtype=torch.float, device=dev)
        self.commands_y = self.commands.view(n, 3)[..., 1]
        self.commands_x = self.commands.view(n, 3)[..., 0]
        self.commands_yaw = self.commands.view(n, 3)[..., 2]

        self.extras = {}
        self.gravity_vec = torch.tensor([0.0, 0.0, -1.0], device=dev).repeat((n, 1))
        self.actions = torch.zeros(n, self.num_actions, dtype=torch.float, device=dev)
        self.last_dof_vel = torch.zeros((n, 12), dtype=torch.float, device=dev)
        self.last_actions = torch.zeros(n, self.num_actions, dtype=torch.float, device=dev)
        self.time_out_buf = torch.zeros_like(self.reset_buf)
        self.fallen_over = torch.zeros(n, dtype=torch.bool, device=dev)
        self.episode_stats = torch.zeros((_lib.STATS_SLOTS, 4), device=dev, dtype=torch.float64)
        self._bind_buffers()

        indices = torch.arange(self._anymals.count, dtype=torch.int64, device=dev)
        self.reset_idx(indices)

    # ---- whole-step entry points --------------------------------------------------------------------------
    def fused_step_enabled(self) -> bool:
        cls = type(self)
        return (not self._hooks_overridden()
                and cls.pre_physics_step is AnymalTask.pre_physics_step
                and cls.post_physics_step is AnymalTask.post_physics_step
                and cls.reset_idx is AnymalTask.reset_idx
                and self._env._world.is_playing())

    def fused_step(self, actions: torch.Tensor, physics_dt: float, graph_mode: bool = False,
                   clip_actions=None) -> None:
        """VecEnvRLGames.step (envs/vec_env_rlgames.py:56-78) as one C-ABI call."""
        self._check_bound()
        a = actions
        if a.dtype != torch.float32 or not a.is_contiguous():
            a = a.float().contiguous()
        self._reset_counter += 1
        rc = _lib.load().oige_anymal_env_step_f32(
            self._anymals.model_handle, C.byref(self._params), C.byref(self._anymals.state_struct),
            C.byref(self._buffers_graph if graph_mode else self._buffers), C.byref(self._reset_init),
            _lib.dptr(a, torch.float32, "actions"), float(self.clip_actions if clip_actions is None else clip_actions), self._reset_seed, self._reset_counter,
            self.env_id_offset, self._num_envs, int(self.control_frequency_inv), float(physics_dt), self._stream())
        _lib.check(rc, "oige_anymal_env_step_f32")

    def read_episode_stats(self, reset: bool = True) -> torch.Tensor:
        """[4] float64 on the device: sum(reward), #episodes finished, sum(length at finish), #env-steps."""
        s = self.episode_stats.sum(dim=0)
        if reset:
            self.episode_stats.zero_()
        return s
