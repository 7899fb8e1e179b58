"""TEST INFRASTRUCTURE ONLY -- drives the VERBATIM reference ``AnymalTask`` methods on the CPU.

``tasks/anymal.py`` and ``robots/articulations/views/anymal_view.py`` are executed unmodified from /root/reference
(never copied) after the Isaac-Sim stubs of ``oracle/ref_loader.py`` (plus the few extra names the two files
import) are installed.  The task is instantiated without its simulator constructor (``object.__new__`` + the
attributes ``AnymalTask.__init__`` / ``get_anymal`` / ``RLTask.__init__`` would have set) on a fake CPU view that
SUBCLASSES the verbatim ``AnymalView`` (so ``is_base_below_threshold`` runs as written); ``post_reset``
(anymal.py:229-250), ``pre_physics_step`` (:181-190), ``reset_idx`` (:192-227), ``get_observations`` (:137-179),
``calculate_metrics`` (:252-287) and ``is_done`` (:289-293) then run exactly as written, in
RLTask.post_physics_step's order (rl_task.py:242-249).  PhysX is unavailable: the state between the pre- and
post-physics calls is injected.  Container-only; writes ``tests/golden/episode_anymal.npz`` via gen_golden.
"""
from typing import Dict, List

import torch

from oracle import anymal_oracle as ao
from oracle import isaac_math as im
from oracle import ref_loader


def load_reference_anymal():
    if not ref_loader.reference_available():
        raise FileNotFoundError(f"reference tree not found at {ref_loader.REFERENCE_ROOT}")
    ref_loader.install_stubs()
    # `from omni.isaac.core.utils.torch.rotations import *` (anymal.py:35): upstream's rotations module itself
    # star-imports the maths module, which is how anymal.py gets torch_rand_float
    ref_loader._mod("omni.isaac.core.utils.torch.rotations", quat_rotate=im.quat_rotate,
                    quat_rotate_inverse=im.quat_rotate_inverse, quat_mul=im.quat_mul,
                    torch_rand_float=im.torch_rand_float, tensor_clamp=im.tensor_clamp)
    ref_loader._mod("omni.isaac.core.prims", RigidPrimView=ref_loader._Dummy)
    ref_loader._mod("omniisaacgymenvs.robots.articulations.anymal", Anymal=ref_loader._Dummy)
    ref_loader._mod("omniisaacgymenvs.robots.articulations.views")
    ref_loader._mod("omniisaacgymenvs.tasks.utils")
    ref_loader._mod("omniisaacgymenvs.tasks.utils.usd_utils", set_drive=lambda *a, **k: None)
    view_mod = ref_loader._load("omniisaacgymenvs.robots.articulations.views.anymal_view",
                                "omniisaacgymenvs/robots/articulations/views/anymal_view.py")
    task_mod = ref_loader._load("omniisaacgymenvs.tasks.anymal", "omniisaacgymenvs/tasks/anymal.py")
    return task_mod, view_mod


def make_fake_view(view_mod, n: int, spawn_pos: torch.Tensor, dof_limits: torch.Tensor):
    class FakeAnymalView(view_mod.AnymalView):  # inherits the verbatim is_base_below_threshold
        def __init__(self):
            self.name, self.count, self.num_dof = "anymalview", n, 12
            self.pos = spawn_pos.clone()
            self.quat = torch.tensor([1.0, 0, 0, 0]).repeat(n, 1)
            self.vel = torch.zeros(n, 6)
            self.dof_pos = torch.zeros(n, 12)
            self.dof_vel = torch.zeros(n, 12)
            self.targets = torch.zeros(n, 12)
            self._limits = dof_limits

        def get_world_poses(self, clone=True):
            return (self.pos.clone(), self.quat.clone()) if clone else (self.pos, self.quat)

        def get_velocities(self, clone=True):
            return self.vel.clone() if clone else self.vel

        def get_joint_positions(self, clone=True):
            return self.dof_pos.clone() if clone else self.dof_pos

        def get_joint_velocities(self, clone=True):
            return self.dof_vel.clone() if clone else self.dof_vel

        def get_dof_limits(self):
            return self._limits.unsqueeze(0).repeat(n, 1, 1)

        def set_joint_position_targets(self, x, indices=None):
            self.targets[indices.long()] = x

        def set_joint_positions(self, x, indices=None):
            self.dof_pos[indices.long()] = x

        def set_joint_velocities(self, x, indices=None):
            self.dof_vel[indices.long()] = x

        def set_world_poses(self, pos, rot, indices=None):
            self.pos[indices.long()] = pos
            self.quat[indices.long()] = rot

        def set_velocities(self, v, indices=None):
            self.vel[indices.long()] = v

    return FakeAnymalView()


def make_reference_task(n: int, spawn_pos: torch.Tensor, dof_limits: torch.Tensor, spec: ao.AnymalSpec = ao.ANYMAL):
    task_mod, view_mod = load_reference_anymal()
    t = object.__new__(task_mod.AnymalTask)
    # AnymalTask.__init__ (anymal.py:54-104)
    t.lin_vel_scale, t.ang_vel_scale = spec.lin_vel_scale, spec.ang_vel_scale
    t.dof_pos_scale, t.dof_vel_scale, t.action_scale = spec.dof_pos_scale, spec.dof_vel_scale, spec.action_scale
    t.rew_scales = {"lin_vel_xy": spec.rew_lin_vel_xy, "ang_vel_z": spec.rew_ang_vel_z, "lin_vel_z": spec.rew_lin_vel_z,
                    "joint_acc": spec.rew_joint_acc, "action_rate": spec.rew_action_rate, "cosmetic": spec.rew_cosmetic}
    t.command_x_range, t.command_y_range = list(spec.command_x_range), list(spec.command_y_range)
    t.command_yaw_range = list(spec.command_yaw_range)
    t.dt = 1 / 60
    t.max_episode_length_s = spec.episode_length_s
    t.max_episode_length = int(t.max_episode_length_s / t.dt + 0.5)
    for key in t.rew_scales.keys():
        t.rew_scales[key] *= t.dt
    t._num_envs, t._device = n, "cpu"
    t._num_observations, t._num_actions = 48, 12
    cls = type(t)
    cls.num_envs = property(lambda self: self._num_envs)
    cls.num_actions = property(lambda self: self._num_actions)
    cls.device = property(lambda self: self._device)
    # get_anymal (anymal.py:129-135)
    t.default_dof_pos = torch.zeros((n, 12), dtype=torch.float)
    for i, name in enumerate(ao.DOF_NAMES):
        t.default_dof_pos[:, i] = spec.default_joint_angles[name]
    # RLTask.__init__ / cleanup (rl_task.py:49-107)
    t.obs_buf = torch.zeros((n, 48))
    t.rew_buf = torch.zeros(n)
    t.reset_buf = torch.ones(n, dtype=torch.long)
    t.progress_buf = torch.zeros(n, dtype=torch.long)
    t.extras = {}
    view = make_fake_view(view_mod, n, spawn_pos, dof_limits)
    t._anymals = view
    return t, view, task_mod


class RandRecorder:
    """Records the raw U[0,1) draws behind the reference module's torch_rand_float calls."""

    def __init__(self, mod):
        self.mod, self.draws = mod, []
        self._orig = mod.torch_rand_float

    def __enter__(self):
        def rec(lower, upper, shape, device):
            u = torch.rand(*shape, device=device)
            self.draws.append(u.clone())
            return (upper - lower) * u + lower

        self.mod.torch_rand_float = rec
        return self

    def __exit__(self, *a):
        self.mod.torch_rand_float = self._orig


def _split(draws: List[torch.Tensor]):
    """One reset_idx call = 4 draws: velocities [k,12], then command x, y, yaw [k,1] each."""
    return draws[0], torch.cat(draws[1:4], dim=1)


def run_episode(n: int, steps: int, seed: int, spawn_pos: torch.Tensor, dof_limits: torch.Tensor,
                spec: ao.AnymalSpec = ao.ANYMAL) -> Dict:
    torch.manual_seed(seed)
    g = torch.Generator().manual_seed(seed + 1000)
    t, view, mod = make_reference_task(n, spawn_pos, dof_limits, spec)
    names = ("reset_mask", "rand_vel", "rand_cmd", "actions", "targets", "commands", "after_reset_dof_pos",
             "after_reset_dof_vel", "after_reset_pos", "after_reset_quat", "after_reset_vel", "after_reset_last_actions",
             "after_reset_last_dof_vel", "inj_pos", "inj_quat", "inj_vel", "inj_dof_pos", "inj_dof_vel", "obs", "rew",
             "reset_buf", "progress_buf", "last_actions", "last_dof_vel")
    rec: Dict[str, List] = {k: [] for k in names}
    with RandRecorder(mod) as rr:
        t.post_reset()
        uv, uc = _split(rr.draws)
        init = {"post_reset_rand_vel": uv, "post_reset_rand_cmd": uc, "post_reset_dof_pos": view.dof_pos.clone(),
                "post_reset_dof_vel": view.dof_vel.clone(), "post_reset_commands": t.commands.clone(),
                "initial_root_pos": t.initial_root_pos.clone(), "initial_root_rot": t.initial_root_rot.clone(),
                "default_dof_pos": t.default_dof_pos.clone(), "dof_lo": t.anymal_dof_lower_limits.clone(),
                "dof_hi": t.anymal_dof_upper_limits.clone()}
        rr.draws.clear()
        for step in range(steps):
            ids = t.reset_buf.nonzero(as_tuple=False).squeeze(-1)
            actions = torch.clamp(torch.rand(n, 12, generator=g) * 2.4 - 1.2, -1.0, 1.0)  # VecEnvRLGames.step clamps first
            t.pre_physics_step(actions)
            mask = torch.zeros(n, dtype=torch.bool)
            mask[ids] = True
            full_v, full_c = torch.zeros(n, 12), torch.zeros(n, 3)
            if len(ids) > 0:
                full_v[ids], full_c[ids] = _split(rr.draws)
            rr.draws.clear()
            for k, v in (("reset_mask", mask), ("rand_vel", full_v), ("rand_cmd", full_c), ("actions", actions),
                         ("targets", view.targets), ("commands", t.commands), ("after_reset_dof_pos", view.dof_pos),
                         ("after_reset_dof_vel", view.dof_vel), ("after_reset_pos", view.pos), ("after_reset_quat", view.quat),
                         ("after_reset_vel", view.vel), ("after_reset_last_actions", t.last_actions),
                         ("after_reset_last_dof_vel", t.last_dof_vel)):
                rec[k].append(v.clone())
            s = ao.synthetic_state(spec, n, seed * 100 + step)
            view.pos[:, :2] = spawn_pos[:, :2] + 0.3 * s["pos"][:, :2]
            view.pos[:, 2] = s["pos"][:, 2]
            view.quat[:] = s["quat"]; view.vel[:] = s["vel"]; view.dof_pos[:] = s["dof_pos"]; view.dof_vel[:] = s["dof_vel"]
            if step == 2:
                t.progress_buf[: n // 4] = t.max_episode_length - 3  # reach the episode limit at the next step
            for k, v in (("inj_pos", view.pos), ("inj_quat", view.quat), ("inj_vel", view.vel), ("inj_dof_pos", view.dof_pos),
                         ("inj_dof_vel", view.dof_vel)):
                rec[k].append(v.clone())
            t.progress_buf[:] += 1
            t.get_observations(); t.calculate_metrics(); t.is_done()
            for k, v in (("obs", t.obs_buf), ("rew", t.rew_buf), ("reset_buf", t.reset_buf), ("progress_buf", t.progress_buf),
                         ("last_actions", t.last_actions), ("last_dof_vel", t.last_dof_vel)):
                rec[k].append(v.clone())
    out = {k: torch.stack(v).numpy() for k, v in rec.items()}
    out.update({k: v.numpy() for k, v in init.items()})
    out["spawn_pos"] = spawn_pos.numpy()
    return out
