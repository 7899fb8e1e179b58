"""TEST INFRASTRUCTURE ONLY -- CPU restatement (torch, fp32) of the reference's flat-terrain Anymal task math.

Follows /root/reference/omniisaacgymenvs/tasks/anymal.py (cited per function) with the third-party helpers of
``oracle/isaac_math.py`` (``quat_rotate``, ``quat_rotate_inverse``, ``torch_rand_float``).  Pinned against the
VERBATIM reference class by ``tests/test_oracle_anymal.py`` (container) and the committed fixture
``tests/golden/episode_anymal.npz`` (``oracle/ref_anymal_driver.py``).  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU legs may import this module.
"""
from dataclasses import dataclass, field
from typing import Dict, List, Tuple

import torch

from oracle import isaac_math as im

DOF_NAMES = ["LF_HAA", "LH_HAA", "RF_HAA", "RH_HAA", "LF_HFE", "LH_HFE", "RF_HFE", "RH_HFE",
             "LF_KFE", "LH_KFE", "RF_KFE", "RH_KFE"]  # robots/articulations/anymal.py:72-83


@dataclass(frozen=True)
class AnymalSpec:
    """cfg/task/Anymal.yaml:6-66 and tasks/anymal.py:54-100."""

    lin_vel_scale: float = 2.0
    ang_vel_scale: float = 0.25
    dof_pos_scale: float = 1.0
    dof_vel_scale: float = 0.05
    action_scale: float = 13.5
    rew_lin_vel_xy: float = 1.0
    rew_ang_vel_z: float = 0.5
    rew_lin_vel_z: float = -0.03
    rew_joint_acc: float = -0.0003
    rew_action_rate: float = -0.006
    rew_cosmetic: float = -0.06
    command_x_range: Tuple[float, float] = (-2.0, 2.0)
    command_y_range: Tuple[float, float] = (-1.0, 1.0)
    command_yaw_range: Tuple[float, float] = (-1.0, 1.0)
    episode_length_s: float = 50.0
    fallen_threshold: float = 0.51          # anymal.py:284
    clip_obs: float = 5.0
    clip_actions: float = 1.0
    num_obs: int = 48
    num_dof: int = 12
    control_frequency_inv: int = 2
    sim_dt: float = 0.01
    spawn_height: float = 0.62
    default_joint_angles: Dict[str, float] = field(default_factory=lambda: {
        "LF_HAA": 0.03, "LH_HAA": 0.03, "RF_HAA": -0.03, "RH_HAA": -0.03,
        "LF_HFE": 0.4, "LH_HFE": -0.4, "RF_HFE": 0.4, "RH_HFE": -0.4,
        "LF_KFE": -0.8, "LH_KFE": 0.8, "RF_KFE": -0.8, "RH_KFE": 0.8})

    @property
    def dt(self) -> float:
        return 1 / 60  # anymal.py:88 (hard-coded; the sim runs 2 x 0.01 s)

    @property
    def max_episode_length(self) -> int:
        return int(self.episode_length_s / self.dt + 0.5)  # anymal.py:90

    def rew_scales(self) -> Dict[str, float]:
        """anymal.py:63-69,94-95: the YAML scales times dt, in Python doubles."""
        s = {"lin_vel_xy": self.rew_lin_vel_xy, "ang_vel_z": self.rew_ang_vel_z, "lin_vel_z": self.rew_lin_vel_z,
             "joint_acc": self.rew_joint_acc, "action_rate": self.rew_action_rate, "cosmetic": self.rew_cosmetic}
        for k in s:
            s[k] *= self.dt
        return s

    def default_dof_pos(self) -> torch.Tensor:
        return torch.tensor([self.default_joint_angles[n] for n in DOF_NAMES], dtype=torch.float32)


ANYMAL = AnymalSpec()


def observe(spec: AnymalSpec, s: Dict[str, torch.Tensor]) -> torch.Tensor:
    """get_observations (anymal.py:137-179), unclamped."""
    velocity, ang_velocity = s["vel"][:, 0:3], s["vel"][:, 3:6]
    base_lin_vel = im.quat_rotate_inverse(s["quat"], velocity) * spec.lin_vel_scale
    base_ang_vel = im.quat_rotate_inverse(s["quat"], ang_velocity) * spec.ang_vel_scale
    gravity_vec = torch.tensor([0.0, 0.0, -1.0]).repeat((s["quat"].shape[0], 1))
    projected_gravity = im.quat_rotate(s["quat"], gravity_vec)
    dof_pos_scaled = (s["dof_pos"] - s["default_dof_pos"]) * spec.dof_pos_scale
    commands_scaled = s["commands"] * torch.tensor([spec.lin_vel_scale, spec.lin_vel_scale, spec.ang_vel_scale])
    return torch.cat((base_lin_vel, base_ang_vel, projected_gravity, commands_scaled, dof_pos_scaled,
                      s["dof_vel"] * spec.dof_vel_scale, s["actions"]), dim=-1)


def metrics(spec: AnymalSpec, s: Dict[str, torch.Tensor]):
    """calculate_metrics (anymal.py:252-287).  Returns (reward, fallen_over) and updates last_actions /
    last_dof_vel in ``s`` like the reference does."""
    rs = spec.rew_scales()
    velocity, ang_velocity = s["vel"][:, 0:3], s["vel"][:, 3:6]
    base_lin_vel = im.quat_rotate_inverse(s["quat"], velocity)
    base_ang_vel = im.quat_rotate_inverse(s["quat"], ang_velocity)
    lin_vel_error = torch.sum(torch.square(s["commands"][:, :2] - base_lin_vel[:, :2]), dim=1)
    ang_vel_error = torch.square(s["commands"][:, 2] - base_ang_vel[:, 2])
    rew_lin_vel_xy = torch.exp(-lin_vel_error / 0.25) * rs["lin_vel_xy"]
    rew_ang_vel_z = torch.exp(-ang_vel_error / 0.25) * rs["ang_vel_z"]
    rew_lin_vel_z = torch.square(base_lin_vel[:, 2]) * rs["lin_vel_z"]
    rew_joint_acc = torch.sum(torch.square(s["last_dof_vel"] - s["dof_vel"]), dim=1) * rs["joint_acc"]
    rew_action_rate = torch.sum(torch.square(s["last_actions"] - s["actions"]), dim=1) * rs["action_rate"]
    rew_cosmetic = torch.sum(torch.abs(s["dof_pos"][:, 0:4] - s["default_dof_pos"][:, 0:4]), dim=1) * rs["cosmetic"]
    total = rew_lin_vel_xy + rew_ang_vel_z + rew_joint_acc + rew_action_rate + rew_cosmetic + rew_lin_vel_z
    total = torch.clip(total, 0.0, None)
    s["last_actions"] = s["actions"].clone()
    s["last_dof_vel"] = s["dof_vel"].clone()
    fallen_over = (s["pos"][:, 2] - 0.0) < spec.fallen_threshold  # AnymalView.is_base_below_threshold
    total[torch.nonzero(fallen_over)] = -1
    return total, fallen_over


def done(spec: AnymalSpec, progress_buf: torch.Tensor, fallen_over: torch.Tensor) -> torch.Tensor:
    """is_done (anymal.py:290-293): bool | bool written into the int64 reset_buf."""
    time_out = progress_buf >= spec.max_episode_length - 1
    return (time_out | fallen_over).long()


def integrate_targets(spec: AnymalSpec, s: Dict[str, torch.Tensor], actions: torch.Tensor) -> None:
    """pre_physics_step after the resets (anymal.py:186-190)."""
    s["actions"] = actions.clone()
    current = s["current_targets"] + spec.action_scale * s["actions"] * spec.dt
    s["current_targets"] = torch.clamp(current, s["dof_lo"], s["dof_hi"])


def reset_idx(spec: AnymalSpec, s: Dict[str, torch.Tensor], ids: torch.Tensor, u_vel: torch.Tensor,
              u_cmd: torch.Tensor) -> None:
    """reset_idx (anymal.py:192-227) from the raw U[0,1) draws: u_vel[k,12] (joint velocities) and u_cmd[k,3]
    (command x, y, yaw, in the reference's draw order)."""
    k = len(ids)
    s["dof_pos"][ids] = s["default_dof_pos"][ids]
    s["dof_vel"][ids] = (0.1 - (-0.1)) * u_vel + (-0.1)
    s["current_targets"][ids] = s["default_dof_pos"][ids]
    s["pos"][ids] = s["initial_root_pos"][ids]
    s["quat"][ids] = s["initial_root_rot"][ids]
    s["vel"][ids] = torch.zeros(k, 6)
    for col, rng in ((0, spec.command_x_range), (1, spec.command_y_range), (2, spec.command_yaw_range)):
        s["commands"][ids, col] = (rng[1] - rng[0]) * u_cmd[:, col] + rng[0]
    s["reset_buf"][ids] = 0
    s["progress_buf"][ids] = 0
    s["last_actions"][ids] = 0.0
    s["last_dof_vel"][ids] = 0.0


def post_physics_step(spec: AnymalSpec, s: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """RLTask.post_physics_step (rl_task.py:242-249): progress += 1, observations, metrics, done."""
    s["progress_buf"] = s["progress_buf"] + 1
    obs = observe(spec, s)
    rew, fallen = metrics(spec, s)
    s["reset_buf"] = done(spec, s["progress_buf"], fallen)
    return {"obs": obs, "rew": rew, "reset_buf": s["reset_buf"], "progress_buf": s["progress_buf"],
            "last_actions": s["last_actions"], "last_dof_vel": s["last_dof_vel"]}


def synthetic_state(spec: AnymalSpec, n: int, seed: int) -> Dict[str, torch.Tensor]:
    """Random post-physics state that straddles the branches: base heights around 0.51, commands in range,
    some rewards clipped at 0, some observations beyond the +-5 clamp."""
    g = torch.Generator().manual_seed(seed)
    r = lambda *shape: torch.randn(*shape, generator=g)
    quat = r(n, 4)
    quat = quat / quat.norm(dim=1, keepdim=True)
    pos = r(n, 3)
    pos[:, 2] = 0.55 + 0.08 * r(n)
    d = spec.default_dof_pos().repeat(n, 1)
    u = torch.rand(n, 3, generator=g)
    cmds = torch.stack([4 * u[:, 0] - 2, 2 * u[:, 1] - 1, 2 * u[:, 2] - 1], dim=1)
    vel = torch.cat([r(n, 3) * 0.8, r(n, 3) * 1.5], dim=1)
    vel[: n // 3, :3] = 0.3 * r(n // 3, 3) + torch.cat([cmds[: n // 3, :2], torch.zeros(n // 3, 1)], dim=1)  # good trackers
    return {"pos": pos.contiguous(), "quat": quat.contiguous(), "vel": vel.contiguous(),
            "dof_pos": (d + 0.4 * r(n, 12)).contiguous(), "dof_vel": (r(n, 12) * 8.0).contiguous(),
            "default_dof_pos": d, "commands": cmds.contiguous(),
            "actions": (torch.rand(n, 12, generator=g) * 2 - 1).contiguous(),
            "last_actions": (torch.rand(n, 12, generator=g) * 2 - 1).contiguous(),
            "last_dof_vel": (r(n, 12) * 8.0).contiguous(),
            "progress_buf": torch.randint(0, 2990, (n,), generator=g),
            "reset_buf": torch.zeros(n, dtype=torch.int64)}
