"""TEST INFRASTRUCTURE ONLY -- drives the VERBATIM reference task METHODS on the CPU.

``oracle/ref_loader.py`` imports the reference's ``tasks/shared/locomotion.py`` / ``tasks/humanoid.py`` /
``tasks/ant.py`` unmodified.  Here the verbatim ``HumanoidLocomotionTask`` / ``AntLocomotionTask``
classes are instantiated WITHOUT their Isaac-Sim constructors (``object.__new__`` + the attributes
``LocomotionTask.__init__`` / ``RLTask.__init__`` would have set) on top of a fake, CPU-tensor
``ArticulationView``, so that the host-side rows of the hot path -- ``post_reset`` (locomotion.py:147-171,
humanoid.py:81-114, ant.py:79-86), ``pre_physics_step`` (:103-114), ``reset_idx`` (:116-145),
``get_observations`` (:80-101), ``calculate_metrics`` (:173-178), ``is_done`` (:180-183) and
``RLTask.post_physics_step``'s ordering (tasks/base/rl_task.py:231-251) -- run exactly as written.
Physics is not available (PhysX): between ``pre_physics_step`` and the post-physics calls the driver
injects synthetic state into the fake view.  Used by ``oracle/gen_golden.py`` to produce the
``episode_*.npz`` fixtures and by CPU tests; needs /root/reference, so container-only.
"""
from typing import Dict, List

import numpy as np
import torch

from oracle import locomotion_oracle as lo
from oracle import ref_loader


class FakePhysicsView:
    def __init__(self, view):
        self._view = view

    def get_force_sensor_forces(self):
        return self._view.sensors


class FakeArticulationView:
    """CPU-tensor stand-in with the 11 methods the tasks call."""

    def __init__(self, n, d, s, spawn_pos, dof_limits):
        self.name, self.count, self.num_dof = "fake_view", n, d
        self.pos = spawn_pos.clone()
        self.quat = torch.tensor([1.0, 0, 0, 0]).repeat(n, 1)
        self.vel = torch.zeros(n, 6)
        self.dof_pos = torch.zeros(n, d)
        self.dof_vel = torch.zeros(n, d)
        self.sensors = torch.zeros(n, s, 6)
        self.efforts = torch.zeros(n, d)
        self._limits = dof_limits  # [D,2]
        self._physics_view = FakePhysicsView(self)

    def get_world_poses(self, clone=True):
        return (self.pos.clone(), self.quat.clone()) if clone else (self.pos, self.quat)

    def get_velocities(self, clone=True):
        return self.vel.clone() if clone else self.vel

    def get_joint_positions(self, clone=True):
        return self.dof_pos.clone() if clone else self.dof_pos

    def get_joint_velocities(self, clone=True):
        return self.dof_vel.clone() if clone else self.dof_vel

    def get_dof_limits(self):
        return self._limits.unsqueeze(0).repeat(self.count, 1, 1)

    def set_joint_efforts(self, forces, indices=None):
        self.efforts[indices.long()] = forces

    def set_joint_positions(self, x, indices=None):
        self.dof_pos[indices.long()] = x

    def set_joint_velocities(self, x, indices=None):
        self.dof_vel[indices.long()] = x

    def set_world_poses(self, pos, rot, indices=None):
        self.pos[indices.long()] = pos
        self.quat[indices.long()] = rot

    def set_velocities(self, v, indices=None):
        self.vel[indices.long()] = v


def make_reference_task(robot: str, n: int, spawn_pos: torch.Tensor, dof_limits: torch.Tensor):
    """Verbatim task object wired to a fake view.  Attribute names are the reference's."""
    loco, hum, ant = ref_loader.load_reference()
    spec = lo.SPECS[robot]
    cls = hum.HumanoidLocomotionTask if robot == "Humanoid" else ant.AntLocomotionTask
    t = object.__new__(cls)
    # what LocomotionTask.__init__ (locomotion.py:53-67) reads from the YAML
    t._num_envs = n
    t._max_episode_length = spec.max_episode_length
    for a in ("dof_vel_scale", "angular_velocity_scale", "contact_force_scale", "power_scale", "heading_weight",
              "up_weight", "actions_cost_scale", "energy_cost_scale", "joints_at_limit_cost_scale", "death_cost",
              "termination_height", "alive_reward_scale"):
        setattr(t, a, getattr(spec, a))
    # what RLTask.__init__ / cleanup (rl_task.py:49-107) set
    t._device = "cpu"
    t._num_observations, t._num_actions = spec.num_obs, spec.num_dof
    type(t).num_envs = property(lambda self: self._num_envs)
    type(t).num_actions = property(lambda self: self._num_actions)
    type(t).num_observations = property(lambda self: self._num_observations)
    t.obs_buf = torch.zeros((n, spec.num_obs))
    t.rew_buf = torch.zeros(n)
    t.reset_buf = torch.ones(n, dtype=torch.long)
    t.progress_buf = torch.zeros(n, dtype=torch.long)
    t.extras = {}
    view = FakeArticulationView(n, spec.num_dof, spec.num_sensors, spawn_pos, dof_limits)
    if robot == "Humanoid":
        t._humanoids = view
    else:
        t._ants = view
    return t, view, loco


class RandRecorder:
    """Wraps the reference module's torch_rand_float so the raw U[0,1) draws are recorded."""

    def __init__(self, loco):
        self.loco, self.draws = loco, []
        self._orig = loco.torch_rand_float

    def __enter__(self):
        def rec(lower, upper, shape, device):
            u = torch.rand(*shape, device=device)
            self.draws.append(u.clone())
            return (upper - lower) * u + lower

        self.loco.torch_rand_float = rec
        return self

    def __exit__(self, *a):
        self.loco.torch_rand_float = self._orig


def run_episode(robot: str, n: int, steps: int, seed: int, dof_limits: torch.Tensor, spawn_pos: torch.Tensor) -> Dict:
    """post_reset, then `steps` x [pre_physics_step -> inject state -> progress+=1 -> get_observations ->
    calculate_metrics -> is_done].  Returns every input and output as numpy arrays."""
    spec = lo.SPECS[robot]
    torch.manual_seed(seed)
    g = torch.Generator().manual_seed(seed + 1000)
    t, view, loco = make_reference_task(robot, n, spawn_pos, dof_limits)
    rec: Dict[str, List] = {k: [] for k in (
        "reset_ids", "rand_pos", "rand_vel", "actions", "efforts", "after_reset_dof_pos", "after_reset_dof_vel",
        "after_reset_pos", "after_reset_quat", "after_reset_vel", "after_reset_potentials", "inj_pos", "inj_quat",
        "inj_vel", "inj_dof_pos", "inj_dof_vel", "inj_sensors", "obs", "rew", "reset_buf", "progress_buf",
        "potentials", "prev_potentials", "up_vec", "heading_vec")}
    with RandRecorder(loco) as rr:
        t.post_reset()
        init = {"initial_root_pos": t.initial_root_pos.clone(), "initial_root_rot": t.initial_root_rot.clone(),
                "initial_dof_pos": t.initial_dof_pos.clone(), "post_reset_rand_pos": rr.draws[0], "post_reset_rand_vel": rr.draws[1],
                "post_reset_dof_pos": view.dof_pos.clone(), "post_reset_dof_vel": view.dof_vel.clone(),
                "post_reset_potentials": t.potentials.clone(), "joint_gears": t.joint_gears.clone(),
                "motor_effort_ratio": t.motor_effort_ratio.clone(), "dof_limits_lower": t.dof_limits_lower.clone(),
                "dof_limits_upper": t.dof_limits_upper.clone()}
        rr.draws.clear()
        for step in range(steps):
            ids = t.reset_buf.nonzero(as_tuple=False).squeeze(-1)
            actions = torch.rand(n, spec.num_dof, generator=g) * 2.4 - 1.2
            actions = torch.clamp(actions, -1.0, 1.0)  # VecEnvRLGames.step clamps first (vec_env_rlgames.py:57)
            t.pre_physics_step(actions)
            mask = torch.zeros(n, dtype=torch.bool)
            mask[ids] = True
            full_rp, full_rv = torch.zeros(n, spec.num_dof), torch.zeros(n, spec.num_dof)
            if len(ids) > 0:
                full_rp[ids], full_rv[ids] = rr.draws[0], rr.draws[1]
            rr.draws.clear()
            rec["reset_ids"].append(mask.clone()); rec["rand_pos"].append(full_rp); rec["rand_vel"].append(full_rv)
            rec["actions"].append(actions.clone()); rec["efforts"].append(view.efforts.clone())
            rec["after_reset_dof_pos"].append(view.dof_pos.clone()); rec["after_reset_dof_vel"].append(view.dof_vel.clone())
            rec["after_reset_pos"].append(view.pos.clone()); rec["after_reset_quat"].append(view.quat.clone())
            rec["after_reset_vel"].append(view.vel.clone()); rec["after_reset_potentials"].append(t.potentials.clone())
            # "physics": inject synthetic state (SURVEY.md 8(d) distributions) around the env origins
            s = lo.synthetic_state(spec, n, seed * 100 + step, dof_lo=t.dof_limits_lower, dof_hi=t.dof_limits_upper)
            view.pos[:, :2] = spawn_pos[:, :2] + 0.3 * s["pos"][:, :2]
            view.pos[:, 2] = s["pos"][:, 2]
            view.quat[:] = s["quat"]; view.vel[:] = s["vel"]; view.dof_pos[:] = s["dof_pos"]; view.dof_vel[:] = s["dof_vel"]
            view.sensors[:] = s["sensors"]
            if step == 2:
                t.progress_buf[: n // 4] = spec.max_episode_length - 2  # some envs hit the episode limit
            for k, v in (("inj_pos", view.pos), ("inj_quat", view.quat), ("inj_vel", view.vel), ("inj_dof_pos", view.dof_pos),
                         ("inj_dof_vel", view.dof_vel), ("inj_sensors", view.sensors)):
                rec[k].append(v.clone())
            # RLTask.post_physics_step (rl_task.py:242-249)
            t.progress_buf[:] += 1
            t.get_observations(); t.calculate_metrics(); t.is_done()
            for k, v in (("obs", t.obs_buf), ("rew", t.rew_buf), ("reset_buf", t.reset_buf), ("progress_buf", t.progress_buf),
                         ("potentials", t.potentials), ("prev_potentials", t.prev_potentials), ("up_vec", t.up_vec),
                         ("heading_vec", t.heading_vec)):
                rec[k].append(v.clone())
    out = {k: torch.stack(v).numpy() for k, v in rec.items()}
    out.update({k: v.numpy() for k, v in init.items()})
    out["spawn_pos"] = spawn_pos.numpy()
    out["progress_override_step"] = np.array(2)
    return out
