"""TEST INFRASTRUCTURE ONLY -- drives the VERBATIM reference ``CartpoleTask`` methods on the CPU.

``tasks/cartpole.py`` is executed unmodified from /root/reference (never copied) after the Isaac-Sim stubs
of ``oracle/ref_loader.py`` are installed.  The class is instantiated without its simulator constructor
(``object.__new__`` + the attributes ``CartpoleTask.__init__`` / ``RLTask.__init__`` would have set) on a
fake CPU ``ArticulationView``; ``post_reset`` (cartpole.py:132-137), ``pre_physics_step`` (:101-108),
``reset_idx`` (:110-130), ``get_observations`` (:80-99), ``calculate_metrics`` (:143-153) and ``is_done``
(:155-162) then run exactly as written, in RLTask.post_physics_step's order (rl_task.py:242-249).  PhysX is
unavailable, so the joint state between the pre- and post-physics calls is injected.  Container-only
(needs /root/reference); ``oracle/gen_golden.py`` uses it to write ``tests/golden/episode_cartpole.npz``.
"""
from typing import Dict, List

import torch

from oracle import cartpole_oracle as co
from oracle import ref_loader


class FakeCartpoleView:
    def __init__(self, n: int):
        self.name, self.count, self.num_dof = "cartpole_view", n, 2
        self.dof_pos = torch.zeros(n, 2)
        self.dof_vel = torch.zeros(n, 2)
        self.efforts = torch.zeros(n, 2)

    def get_dof_index(self, name: str) -> int:
        return {"cartJoint": 0, "poleJoint": 1}[name]

    def get_joint_positions(self, clone=True):
        return self.dof_pos.clone() if clone else self.dof_pos

    def get_joint_velocities(self, clone=True):
        return self.dof_vel.clone() if clone else self.dof_vel

    def set_joint_efforts(self, forces, indices=None):
        self.efforts[indices.long()] = forces

    def set_joint_positions(self, x, indices=None):
        self.dof_pos[indices.long()] = x

    def set_joint_velocities(self, x, indices=None):
        self.dof_vel[indices.long()] = x


def load_reference_cartpole():
    if not ref_loader.reference_available():
        raise FileNotFoundError(f"reference tree not found at {ref_loader.REFERENCE_ROOT}")
    ref_loader.install_stubs()
    ref_loader._mod("omniisaacgymenvs.robots.articulations.cartpole", Cartpole=ref_loader._Dummy)
    return ref_loader._load("omniisaacgymenvs.tasks.cartpole", "omniisaacgymenvs/tasks/cartpole.py")


def make_reference_task(n: int, spec: co.CartpoleSpec = co.CARTPOLE):
    mod = load_reference_cartpole()
    t = object.__new__(mod.CartpoleTask)
    # CartpoleTask.__init__ (cartpole.py:53-66) / RLTask.__init__ + cleanup (rl_task.py:49-107)
    t._num_envs, t._device = n, "cpu"
    t._reset_dist, t._max_push_effort, t._max_episode_length = spec.reset_dist, spec.max_push_effort, 500
    t._num_observations, t._num_actions = 4, 1
    t.obs_buf = torch.zeros((n, 4))
    t.rew_buf = torch.zeros(n)
    t.reset_buf = torch.ones(n, dtype=torch.long)
    t.progress_buf = torch.zeros(n, dtype=torch.long)
    t.extras = {}
    view = FakeCartpoleView(n)
    t._cartpoles = view
    return t, view, mod


class RandRecorder:
    """Records the raw ``torch.rand`` draws the reference makes in reset_idx (four per call, cartpole.py:115-121)."""

    def __init__(self):
        self.draws: List[torch.Tensor] = []
        self._orig = torch.rand

    def __enter__(self):
        def rec(*a, **k):
            u = self._orig(*a, **k)
            self.draws.append(u.clone())
            return u

        torch.rand = rec
        return self

    def __exit__(self, *a):
        torch.rand = self._orig


def run_episode(n: int, steps: int, seed: int, spec: co.CartpoleSpec = co.CARTPOLE) -> Dict:
    torch.manual_seed(seed)
    g = torch.Generator().manual_seed(seed + 1000)
    t, view, _ = make_reference_task(n, spec)
    rec: Dict[str, List] = {k: [] for k in ("reset_mask", "rand", "actions", "efforts", "after_reset_dof_pos",
                                            "after_reset_dof_vel", "inj_dof_pos", "inj_dof_vel", "obs", "rew",
                                            "reset_buf", "progress_buf")}
    with RandRecorder() as rr:
        t.post_reset()
        init = {"post_reset_rand": torch.stack(rr.draws, dim=1), "post_reset_dof_pos": view.dof_pos.clone(),
                "post_reset_dof_vel": view.dof_vel.clone()}
        rr.draws.clear()
        for step in range(steps):
            ids = t.reset_buf.nonzero(as_tuple=False).squeeze(-1)
            # the random policy draws through the unpatched torch.rand; VecEnvRLGames.step clamps first
            actions = torch.clamp(rr._orig(n, 1, generator=g) * 2.4 - 1.2, -1.0, 1.0)
            t.pre_physics_step(actions)
            mask = torch.zeros(n, dtype=torch.bool)
            mask[ids] = True
            full = torch.zeros(n, 4)
            if len(ids) > 0:
                full[ids] = torch.stack(rr.draws, dim=1)
            rr.draws.clear()
            rec["reset_mask"].append(mask); rec["rand"].append(full); rec["actions"].append(actions.clone())
            rec["efforts"].append(view.efforts.clone())
            rec["after_reset_dof_pos"].append(view.dof_pos.clone()); rec["after_reset_dof_vel"].append(view.dof_vel.clone())
            q, qd = co.synthetic_state(n, seed * 100 + step)
            view.dof_pos[:], view.dof_vel[:] = q, qd
            if step == 2:
                t.progress_buf[: n // 4] = 498  # reach the episode limit at this and the next step
            rec["inj_dof_pos"].append(q.clone()); rec["inj_dof_vel"].append(qd.clone())
            t.progress_buf[:] += 1
            t.get_observations(); t.calculate_metrics(); t.is_done()
            for k, v in (("obs", t.obs_buf), ("rew", t.rew_buf), ("reset_buf", t.reset_buf), ("progress_buf", t.progress_buf)):
                rec[k].append(v.clone())
    out = {k: torch.stack(v).numpy() for k, v in rec.items()}
    out.update({k: v.numpy() for k, v in init.items()})
    return out
