"""TEST INFRASTRUCTURE ONLY -- CPU oracle for the obs / reward / termination / reset pass.

A torch-fp32 restatement of the reference's locomotion task arithmetic, operation for
operation in the reference's order so that every non-transcendental output is bit-identical
to the verbatim reference functions (pinned by ``tests/test_oracle_golden.py`` here in
the container, and by the committed fixtures under ``tests/golden/`` everywhere else).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl
reference`` legs may import this module.  The product package never does.

PARITY NOTE: the reference ships no tests/golden vectors; the pin is the
reference's own functions executed in this container via ``oracle/ref_loader.py`` -> fixtures.
The third-party helpers (``oracle/isaac_math.py``) are "parity unpinned" (source absent).

Functions and the reference lines they follow (paths under /root/reference/omniisaacgymenvs):
  observe()            tasks/shared/locomotion.py:194-254 (jit get_observations), :190-192
  limit_cost_humanoid  tasks/humanoid.py:120-127
  limit_cost_ant       tasks/ant.py:92-95
  reward()             tasks/shared/locomotion.py:271-321 (jit calculate_metrics)
  done()               tasks/shared/locomotion.py:257-268 (jit is_done)
  post_physics_step()  tasks/base/rl_task.py:231-251 + envs/vec_env_rlgames.py:41-46 (obs clamp)
  reset_idx()          tasks/shared/locomotion.py:116-145
  joint_efforts()      tasks/shared/locomotion.py:108-109
"""
from dataclasses import dataclass, field
from typing import Dict, List

import torch

from oracle import isaac_math as im

HUMANOID_GEARS = [67.5] * 7 + [45.0, 45.0] + [45.0, 135.0, 45.0] * 2 + [90.0, 90.0] + [22.5] * 4  # tasks/humanoid.py:82-107


@dataclass
class TaskSpec:
    """Scalar constants of one locomotion task (cfg/task/{Humanoid,Ant}.yaml:9-32)."""

    name: str
    num_dof: int
    num_sensors: int
    num_obs: int
    dof_vel_scale: float
    angular_velocity_scale: float
    contact_force_scale: float
    power_scale: float
    heading_weight: float
    up_weight: float
    actions_cost_scale: float
    energy_cost_scale: float
    joints_at_limit_cost_scale: float
    death_cost: float
    termination_height: float
    alive_reward_scale: float
    max_episode_length: int = 1000
    dt: float = 1.0 / 60.0  # tasks/shared/locomotion.py:163 (hard-coded, not sim dt)
    clip_obs: float = float("inf")  # tasks/base/rl_task.py:69
    joint_gears: List[float] = field(default_factory=list)

    @property
    def limit_cost_kind(self) -> str:
        return "ant" if self.name == "Ant" else "humanoid"

    def motor_effort_ratio(self) -> torch.Tensor:
        g = torch.tensor(self.joint_gears, dtype=torch.float32)
        if self.name == "Ant":  # tasks/ant.py:84
            return torch.ones_like(g)
        return g / torch.max(g)  # tasks/humanoid.py:108-109


HUMANOID = TaskSpec(
    name="Humanoid", num_dof=21, num_sensors=2, num_obs=87,
    dof_vel_scale=0.1, angular_velocity_scale=0.25, contact_force_scale=0.01, power_scale=1.0,
    heading_weight=0.5, up_weight=0.1, actions_cost_scale=0.01, energy_cost_scale=0.05,
    joints_at_limit_cost_scale=0.25, death_cost=-1.0, termination_height=0.8, alive_reward_scale=2.0,
    joint_gears=HUMANOID_GEARS,
)
ANT = TaskSpec(
    name="Ant", num_dof=8, num_sensors=4, num_obs=60,
    dof_vel_scale=0.2, angular_velocity_scale=1.0, contact_force_scale=0.1, power_scale=0.5,
    heading_weight=0.5, up_weight=0.1, actions_cost_scale=0.005, energy_cost_scale=0.05,
    joints_at_limit_cost_scale=0.1, death_cost=-2.0, termination_height=0.31, alive_reward_scale=0.5,
    joint_gears=[15.0] * 8,
)
SPECS = {"Humanoid": HUMANOID, "Ant": ANT}


def normalize_angle(x):
    return torch.atan2(torch.sin(x), torch.cos(x))


def observe(spec: TaskSpec, s: Dict[str, torch.Tensor]):
    """s holds: pos[N,3] quat[N,4] vel[N,6] dof_pos[N,D] dof_vel[N,D] sensors[N,S,6] actions[N,D]
    potentials[N] targets[N,3] inv_start_rot[N,4] basis_vec0[N,3] basis_vec1[N,3] dof_lo[D] dof_hi[D].
    Returns (obs, potentials, prev_potentials, up_vec, heading_vec)."""
    n = s["pos"].shape[0]
    pos, quat = s["pos"], s["quat"]
    velocity, ang_velocity = s["vel"][:, 0:3], s["vel"][:, 3:6]
    to_target = s["targets"] - pos
    to_target[:, 2] = 0.0
    prev_potentials = s["potentials"].clone()
    potentials = -torch.norm(to_target, p=2, dim=-1) / spec.dt
    torso_quat, up_proj, heading_proj, up_vec, heading_vec = im.compute_heading_and_up(
        quat, s["inv_start_rot"], to_target, s["basis_vec0"], s["basis_vec1"], 2
    )
    vel_loc, angvel_loc, roll, _pitch, yaw, angle_to_target = im.compute_rot(
        torso_quat, velocity, ang_velocity, s["targets"], pos
    )
    dof_pos_scaled = im.unscale(s["dof_pos"], s["dof_lo"], s["dof_hi"])
    obs = torch.cat(
        (
            pos[:, 2].view(-1, 1),
            vel_loc,
            angvel_loc * spec.angular_velocity_scale,
            normalize_angle(yaw).unsqueeze(-1),
            normalize_angle(roll).unsqueeze(-1),
            normalize_angle(angle_to_target).unsqueeze(-1),
            up_proj.unsqueeze(-1),
            heading_proj.unsqueeze(-1),
            dof_pos_scaled,
            s["dof_vel"] * spec.dof_vel_scale,
            s["sensors"].reshape(n, -1) * spec.contact_force_scale,
            s["actions"],
        ),
        dim=-1,
    )
    return obs, potentials, prev_potentials, up_vec, heading_vec


def limit_cost_humanoid(spec: TaskSpec, obs, motor_effort_ratio):
    d = spec.num_dof
    scaled_cost = spec.joints_at_limit_cost_scale * (torch.abs(obs[:, 12:12 + d]) - 0.98) / 0.02
    return torch.sum((torch.abs(obs[:, 12:12 + d]) > 0.98) * scaled_cost * motor_effort_ratio.unsqueeze(0), dim=-1)


def limit_cost_ant(spec: TaskSpec, obs):
    return torch.sum(obs[:, 12:12 + spec.num_dof] > 0.99, dim=-1)  # int64, no abs, unscaled


def limit_cost(spec: TaskSpec, obs, motor_effort_ratio):
    if spec.limit_cost_kind == "ant":
        return limit_cost_ant(spec, obs)
    return limit_cost_humanoid(spec, obs, motor_effort_ratio)


def reward(spec: TaskSpec, obs, actions, potentials, prev_potentials, dof_at_limit_cost, motor_effort_ratio):
    d = spec.num_dof
    heading_weight_tensor = torch.ones_like(obs[:, 11]) * spec.heading_weight
    heading_reward = torch.where(obs[:, 11] > 0.8, heading_weight_tensor, spec.heading_weight * obs[:, 11] / 0.8)
    up_reward = torch.zeros_like(heading_reward)
    up_reward = torch.where(obs[:, 10] > 0.93, up_reward + spec.up_weight, up_reward)
    actions_cost = torch.sum(actions ** 2, dim=-1)
    electricity_cost = torch.sum(
        torch.abs(actions * obs[:, 12 + d:12 + d * 2]) * motor_effort_ratio.unsqueeze(0), dim=-1
    )
    alive_reward = torch.ones_like(potentials) * spec.alive_reward_scale
    progress_reward = potentials - prev_potentials
    total = (
        progress_reward
        + alive_reward
        + up_reward
        + heading_reward
        - spec.actions_cost_scale * actions_cost
        - spec.energy_cost_scale * electricity_cost
        - dof_at_limit_cost
    )
    total = torch.where(obs[:, 0] < spec.termination_height, torch.ones_like(total) * spec.death_cost, total)
    return total


def done(spec: TaskSpec, obs, reset_buf, progress_buf):
    max_len = float(spec.max_episode_length)  # the jit signature takes a float (locomotion.py:265)
    r = torch.where(obs[:, 0] < spec.termination_height, torch.ones_like(reset_buf), reset_buf)
    r = torch.where(progress_buf >= max_len - 1, torch.ones_like(reset_buf), r)
    return r


def post_physics_step(spec: TaskSpec, s: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """progress += 1 -> observe -> limit cost -> reward -> done -> obs clamp.  Pure: returns new tensors."""
    mer = spec.motor_effort_ratio()
    progress = s["progress_buf"] + 1
    obs, pot, prev_pot, up_vec, heading_vec = observe(spec, s)
    cost = limit_cost(spec, obs, mer)
    rew = reward(spec, obs, s["actions"], pot, prev_pot, cost, mer)
    reset = done(spec, obs, s["reset_buf"], progress)
    obs_out = torch.clamp(obs, -spec.clip_obs, spec.clip_obs)
    return {
        "obs": obs_out, "rew": rew, "reset_buf": reset, "progress_buf": progress,
        "potentials": pot, "prev_potentials": prev_pot, "up_vec": up_vec, "heading_vec": heading_vec,
    }


def joint_efforts(spec: TaskSpec, actions):
    gears = torch.tensor(spec.joint_gears, dtype=torch.float32)
    return actions * gears * spec.power_scale


def reset_idx(spec: TaskSpec, s: Dict[str, torch.Tensor], env_ids, rand_pos, rand_vel):
    """In-place reset of rows env_ids.  rand_pos / rand_vel are the raw U[0,1) draws [k,D] that
    torch_rand_float would have made (locomotion.py:120,124), so the arithmetic is reproducible."""
    dof_pos = (0.2 - (-0.2)) * rand_pos + (-0.2)
    dof_pos = im.tensor_clamp(s["initial_dof_pos"][env_ids] + dof_pos, s["dof_lo"], s["dof_hi"])
    dof_vel = (0.1 - (-0.1)) * rand_vel + (-0.1)
    s["dof_pos"][env_ids] = dof_pos
    s["dof_vel"][env_ids] = dof_vel
    s["pos"][env_ids] = s["initial_root_pos"][env_ids]
    s["quat"][env_ids] = s["initial_root_rot"][env_ids]
    s["vel"][env_ids] = 0.0
    to_target = s["targets"][env_ids] - s["initial_root_pos"][env_ids]
    to_target[:, 2] = 0.0
    s["prev_potentials"][env_ids] = -torch.norm(to_target, p=2, dim=-1) / spec.dt
    s["potentials"][env_ids] = s["prev_potentials"][env_ids].clone()
    s["reset_buf"][env_ids] = 0
    s["progress_buf"][env_ids] = 0


def synthetic_state(spec: TaskSpec, n: int, seed: int, dof_lo=None, dof_hi=None) -> Dict[str, torch.Tensor]:
    """Synthetic K1 inputs of SURVEY.md section 8(d) (CPU generator, torch.manual_seed(seed))."""
    g = torch.Generator().manual_seed(seed)
    d, ns = spec.num_dof, spec.num_sensors
    humanoid = spec.name != "Ant"
    pos = torch.randn(n, 3, generator=g)
    pos[:, 2] = (1.0 + 0.3 * torch.randn(n, generator=g)) if humanoid else (0.45 + 0.15 * torch.randn(n, generator=g))
    quat = torch.randn(n, 4, generator=g)
    quat = quat / quat.norm(dim=-1, keepdim=True)
    vel = torch.randn(n, 6, generator=g)
    if dof_lo is None:
        dof_lo = -(0.5 + torch.rand(d, generator=g))
        dof_hi = 0.5 + torch.rand(d, generator=g)
    rng = dof_hi - dof_lo
    u = torch.rand(n, d, generator=g)
    dof_pos = (dof_lo - 0.05 * rng) + u * (1.1 * rng)
    dof_vel = torch.randn(n, d, generator=g)
    sensors = torch.randn(n, ns, 6, generator=g) * 10.0
    actions = torch.rand(n, d, generator=g) * 2.0 - 1.0
    potentials = -1000.0 * 60.0 + torch.randn(n, generator=g)
    progress = torch.randint(0, 1000, (n,), generator=g, dtype=torch.int64)
    force = torch.rand(n, generator=g) < 0.01
    progress = torch.where(force, torch.randint(997, 1000, (n,), generator=g, dtype=torch.int64), progress)
    reset = torch.zeros(n, dtype=torch.int64)
    one = lambda v: torch.tensor(v, dtype=torch.float32).repeat(n, 1)
    return {
        "pos": pos, "quat": quat, "vel": vel, "dof_pos": dof_pos, "dof_vel": dof_vel, "sensors": sensors,
        "actions": actions, "potentials": potentials, "progress_buf": progress, "reset_buf": reset,
        "targets": one([1000.0, 0.0, 0.0]), "inv_start_rot": one([1.0, 0.0, 0.0, 0.0]),
        "basis_vec0": one([1.0, 0.0, 0.0]), "basis_vec1": one([0.0, 0.0, 1.0]),
        "dof_lo": dof_lo, "dof_hi": dof_hi,
    }
