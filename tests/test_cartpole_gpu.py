"""GPU: the Cartpole passes through the C ABI against oracle/cartpole_oracle.py (pinned bit-exact to the verbatim
reference CartpoleTask, tests/test_oracle_cartpole.py) and the task / env classes end to end.

Bar: every output is BIT-exact (obs, rew, reset_buf, progress_buf, reset values, efforts) -- the passes are
elementwise fp32 with no reductions.  Physics (K2 on the cart-pole tree) is covered by tests/test_k2_dynamics_gpu.py."""
import ctypes as C
import os

import numpy as np
import pytest
import torch

import helpers
from omniisaacgymenvs_b200 import _lib
from omniisaacgymenvs_b200.utils.task_util import make_env
from oracle import cartpole_oracle as co
from oracle import philox

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _params(inc=1, cart=0, clip=5.0):
    p = _lib.OigeCartpoleParams()
    p.cart_dof, p.pole_dof, p.progress_increment = cart, 1 - cart, inc
    p.reset_dist, p.max_push_effort, p.max_episode_length, p.clip_obs = 3.0, 400.0, 500.0, clip
    return p


class Case:
    def __init__(self, n, q=None, qd=None, progress=None, reset=None):
        z = lambda *s, dt=torch.float32: torch.zeros(*s, dtype=dt, device=DEV)
        self.n = n
        self.q = z(n, 2) if q is None else q.to(DEV).contiguous()
        self.qd = z(n, 2) if qd is None else qd.to(DEV).contiguous()
        self.obs, self.rew, self.eff = z(n, 4), z(n), z(n, 2)
        self.reset = z(n, dt=torch.int64) if reset is None else reset.to(DEV).contiguous()
        self.progress = z(n, dt=torch.int64) if progress is None else progress.to(DEV).contiguous()
        self.stats = z(_lib.STATS_SLOTS, 4, dt=torch.float64)
        self.state = _lib.OigeState(None, None, None, self.q.data_ptr(), self.qd.data_ptr(), None, None)
        self.buf = _lib.OigeCartpoleBuffers(self.obs.data_ptr(), self.rew.data_ptr(), self.reset.data_ptr(),
                                            self.progress.data_ptr(), self.eff.data_ptr(), self.stats.data_ptr(), None)

    def post(self, lib, p):
        _lib.check(lib.oige_cartpole_obs_reward_reset_f32(C.byref(p), C.byref(self.state), C.byref(self.buf), self.n,
                                                          _lib.current_stream_ptr(DEV)), "post")
        torch.cuda.synchronize()


@pytest.mark.parametrize("n", [1, 33, 4096, 100001])
@pytest.mark.parametrize("cart", [0, 1])
def test_obs_reward_done_bit_exact(cuda_lib, n, cart):
    spec = co.CartpoleSpec(cart_dof=cart, pole_dof=1 - cart)
    q, qd = co.synthetic_state(n, 5)
    if cart == 1:
        q, qd = q.flip(1).contiguous(), qd.flip(1).contiguous()
    g = torch.Generator().manual_seed(1)
    progress = torch.randint(0, 497, (n,), generator=g)
    progress[::7] = 499  # -> 500 after the increment: the episode limit
    progress[1::7] = 498
    c = Case(n, q, qd, progress, torch.ones(n, dtype=torch.int64))  # incoming reset flags are NOT carried over
    c.post(cuda_lib, _params(cart=cart))
    obs, rew, reset, prog = co.post_physics_step(spec, q, qd, progress)
    assert torch.equal(c.obs.cpu(), torch.clamp(obs, -5.0, 5.0))
    assert torch.equal(c.rew.cpu(), rew)
    assert torch.equal(c.reset.cpu(), reset) and torch.equal(c.progress.cpu(), prog)
    st = c.stats.sum(0).cpu()
    assert int(st[3]) == n and int(st[1]) == int(reset.sum())
    assert abs(float(st[0]) - float(rew.double().sum())) < 1e-6 * max(1.0, n)
    assert int(st[2]) == int(prog[reset != 0].sum())


def test_progress_increment_off_and_no_clip(cuda_lib):
    q, qd = co.synthetic_state(64, 2)
    progress = torch.full((64,), 499, dtype=torch.int64)
    c = Case(64, q, qd, progress)
    c.post(cuda_lib, _params(inc=0, clip=float("inf")))
    obs = co.observe(co.CARTPOLE, q, qd)
    assert torch.equal(c.obs.cpu(), obs) and torch.equal(c.progress.cpu(), progress)
    assert torch.equal(c.reset.cpu(), co.done(co.CARTPOLE, obs, progress))


def test_reference_episode_through_the_c_abi(cuda_lib):
    """tests/golden/episode_cartpole.npz (recorded from the verbatim reference class) replayed kernel by kernel."""
    z = np.load(os.path.join(helpers.GOLDEN_DIR, "episode_cartpole.npz"))
    ep = {k: torch.from_numpy(z[k]) for k in z.files}
    steps, n = ep["actions"].shape[:2]
    lib, p, s = cuda_lib, _params(), _lib.current_stream_ptr(DEV)
    c = Case(n, reset=torch.ones(n, dtype=torch.int64))
    ids = torch.arange(n, device=DEV)
    rand = ep["post_reset_rand"].to(DEV).contiguous()
    _lib.check(lib.oige_cartpole_reset_idx_f32(C.byref(p), C.byref(c.state), C.byref(c.buf), ids.data_ptr(), n,
                                               rand.data_ptr(), 0, 0, 0, s), "reset_idx")
    assert torch.equal(c.q.cpu(), ep["post_reset_dof_pos"]) and torch.equal(c.qd.cpu(), ep["post_reset_dof_vel"])
    assert int(c.reset.sum()) == 0
    for t in range(steps):
        assert torch.equal(c.reset.cpu() != 0, ep["reset_mask"][t]), f"step {t}: reset set"
        a = ep["actions"][t].to(DEV).contiguous()
        rand = ep["rand"][t].to(DEV).contiguous()
        _lib.check(lib.oige_cartpole_reset_masked_f32(C.byref(p), C.byref(c.state), C.byref(c.buf), a.data_ptr(), 1.0,
                                                      rand.data_ptr(), 0, 0, 0, n, s), "pre")
        assert torch.equal(c.q.cpu(), ep["after_reset_dof_pos"][t]) and torch.equal(c.qd.cpu(), ep["after_reset_dof_vel"][t])
        assert torch.equal(c.eff.cpu(), ep["efforts"][t]), f"step {t}: efforts"
        assert int(c.reset.sum()) == 0
        c.q.copy_(ep["inj_dof_pos"][t]); c.qd.copy_(ep["inj_dof_vel"][t])
        if t == 2:
            c.progress[: n // 4] = 498
        c.post(lib, p)
        assert torch.equal(c.obs.cpu(), torch.clamp(ep["obs"][t], -5.0, 5.0)), f"step {t}: obs"
        assert torch.equal(c.rew.cpu(), ep["rew"][t]), f"step {t}: rew"
        assert torch.equal(c.reset.cpu(), ep["reset_buf"][t]) and torch.equal(c.progress.cpu(), ep["progress_buf"][t])


def test_philox_reset_is_bit_exact_and_sharding_invariant(cuda_lib):
    n, seed, counter, off = 1000, 0x1234ABCD5678, 77, 5000
    reset = (torch.arange(n) % 3 != 0).long()
    q0 = torch.full((n, 2), 9.0)
    c = Case(n, q0.clone(), q0.clone(), torch.full((n,), 5, dtype=torch.int64), reset)
    a = (torch.rand(n, 1) * 3 - 1.5).to(DEV)
    _lib.check(cuda_lib.oige_cartpole_reset_masked_f32(C.byref(_params()), C.byref(c.state), C.byref(c.buf), a.data_ptr(),
                                                       1.0, None, seed, counter, off, n, _lib.current_stream_ptr(DEV)), "pre")
    u = torch.from_numpy(philox.reset_uniforms(seed, counter, np.arange(off, off + n), 4, 0))
    eq, eqd = co.reset_values(u)
    m = reset.bool()
    assert torch.equal(c.q.cpu()[m], eq[m]) and torch.equal(c.qd.cpu()[m], eqd[m])
    assert torch.equal(c.q.cpu()[~m], q0[~m]) and torch.equal(c.progress.cpu()[~m], torch.full((int((~m).sum()),), 5))
    assert int(c.reset.sum()) == 0 and int(c.progress.cpu()[m].sum()) == 0
    assert torch.equal(c.eff.cpu(), co.efforts(co.CARTPOLE, torch.clamp(a.cpu(), -1.0, 1.0)))
    # a shard that owns global envs [off+400, off+700) draws the same numbers
    c2 = Case(300, reset=torch.ones(300, dtype=torch.int64))
    _lib.check(cuda_lib.oige_cartpole_reset_masked_f32(C.byref(_params()), C.byref(c2.state), C.byref(c2.buf), None, 1.0,
                                                       None, seed, counter, off + 400, 300, _lib.current_stream_ptr(DEV)), "pre")
    assert torch.equal(c2.q.cpu(), eq[400:700]) and torch.equal(c2.qd.cpu(), eqd[400:700])
    # reset noise ranges (cartpole.py:115-121)
    assert float(eq[:, 0].abs().max()) <= 1.0 and float(eq[:, 1].abs().max()) <= 0.125 * np.pi + 1e-6


def test_argument_errors_come_back_as_codes(cuda_lib):
    c = Case(8)
    p = _params()
    p.pole_dof = 0  # not a permutation
    assert cuda_lib.oige_cartpole_obs_reward_reset_f32(C.byref(p), C.byref(c.state), C.byref(c.buf), 8, None) == -2
    assert b"permutation" in cuda_lib.oige_last_error()
    assert cuda_lib.oige_cartpole_obs_reward_reset_f32(None, C.byref(c.state), C.byref(c.buf), 8, None) == -1
    assert cuda_lib.oige_cartpole_env_step_f32(None, C.byref(_params()), C.byref(c.state), C.byref(c.buf), None, 1.0, 0, 0,
                                               0, 8, 2, 0.0083, None) == -1


# ---- task / env level ---------------------------------------------------------------------------------------
def _snap(task):
    v = task.get_robot()
    return {k: t.clone() for k, t in (("q", v.dof_pos), ("qd", v.dof_vel), ("obs", task.obs_buf), ("rew", task.rew_buf),
                                      ("reset", task.reset_buf), ("progress", task.progress_buf), ("eff", v.dof_efforts))}


def test_env_surface_matches_the_reference_task(cuda_lib):
    env, task = make_env("Cartpole", num_envs=512, seed=0)
    assert task.num_observations == 4 and task.num_actions == 1 and task._max_episode_length == 500
    assert task.clip_obs == 5.0 and task.clip_actions == 1.0 and task.control_frequency_inv == 2
    v = task.get_robot()
    assert v.num_dof == 2 and v.get_dof_index("cartJoint") == 0 and v.get_dof_index("poleJoint") == 1
    assert float(v.root_pos[:, 2].min()) == 2.0  # cartpole.py:57
    obs = env.reset()
    assert obs["obs"].shape == (512, 4) and float(obs["obs"].abs().max()) <= 5.0
    o, r, d, extras = env.step(torch.zeros(512, 1, device=task.device))
    assert o["obs"].shape == (512, 4) and r.shape == (512,) and d.dtype == torch.int64 and extras == {}
    # post_reset drew the start state inside the reference's ranges (cartpole.py:115-121)
    env2, task2 = make_env("Cartpole", num_envs=4096, seed=1)
    q, qd = task2.get_robot().dof_pos, task2.get_robot().dof_vel
    assert float(q[:, 0].abs().max()) <= 1.0 and float(q[:, 1].abs().max()) <= 0.125 * np.pi + 1e-6
    assert float(qd[:, 0].abs().max()) <= 0.5 and float(qd[:, 1].abs().max()) <= 0.25 * np.pi + 1e-6
    assert float(q[:, 0].std()) > 0.5 and abs(float(q[:, 0].mean())) < 0.05


def test_fused_step_equals_hook_by_hook_path(cuda_lib):
    """One C-ABI call (3 launches) vs the reference's call order through the overridable hooks."""
    from omniisaacgymenvs_b200.tasks.cartpole import CartpoleTask

    class Hooked(CartpoleTask):  # overriding a hook turns the fused path off
        def is_done(self):
            CartpoleTask.is_done(self)

    env_a, task_a = make_env("Cartpole", num_envs=700, seed=3)
    env_b, task_b = make_env("Cartpole", num_envs=700, seed=3)
    task_b.__class__ = Hooked
    assert task_a.fused_step_enabled() and not task_b.fused_step_enabled()
    env_a.reset(); env_b.reset()
    g = torch.Generator().manual_seed(0)
    resets = 0
    for step in range(120):
        a = (torch.rand(700, 1, generator=g) * 2.4 - 1.2).to(task_a.device)
        oa, ra, da, _ = env_a.step(a)
        ob, rb, db, _ = env_b.step(a)
        assert torch.equal(oa["obs"], ob["obs"]), f"step {step}"
        assert torch.equal(ra, rb) and torch.equal(da, db)
        resets += int(da.sum())
    assert resets > 100
    sa, sb = _snap(task_a), _snap(task_b)
    for k in sa:
        assert torch.equal(sa[k], sb[k]), k


def test_post_physics_step_matches_oracle_on_live_state(cuda_lib):
    env, task = make_env("Cartpole", num_envs=2048, seed=5)
    env.reset()
    v = task.get_robot()
    for step in range(40):
        a = torch.rand(2048, 1, device=task.device) * 2 - 1
        prog = task.progress_buf.clone()
        flagged = task.reset_buf.bool().cpu()
        o, r, d, _ = env.step(a)
        prog = torch.where(flagged.to(prog.device), torch.zeros_like(prog), prog)
        obs, rew, reset, prog2 = co.post_physics_step(co.CARTPOLE, v.dof_pos.cpu(), v.dof_vel.cpu(), prog.cpu())
        assert torch.equal(o["obs"].cpu(), torch.clamp(obs, -5.0, 5.0)) and torch.equal(r.cpu(), rew)
        assert torch.equal(d.cpu(), reset) and torch.equal(task.progress_buf.cpu(), prog2)
        assert torch.equal(v.dof_efforts.cpu(), co.efforts(co.CARTPOLE, a.cpu()))


def test_sharded_envs_equal_rows_of_the_unsharded_run(cuda_lib):
    n = 512
    env_full, task_full = make_env("Cartpole", num_envs=n, seed=11)
    env_half, task_half = make_env("Cartpole", num_envs=n // 2, seed=11, env_id_offset=n // 2, global_num_envs=n)
    env_full.reset(); env_half.reset()
    g = torch.Generator().manual_seed(2)
    for step in range(80):
        a = (torch.rand(n, 1, generator=g) * 2 - 1).to(task_full.device)
        of, rf, df, _ = env_full.step(a)
        oh, rh, dh_, _ = env_half.step(a[n // 2:].contiguous())
        assert torch.equal(of["obs"][n // 2:], oh["obs"]), f"step {step}"
        assert torch.equal(rf[n // 2:], rh) and torch.equal(df[n // 2:], dh_)
    assert int(task_full.read_episode_stats(reset=False)[3]) == n * 81


@pytest.mark.parametrize("n", [512, 16384])
def test_cuda_graph_step_equals_eager_step(cuda_lib, n):
    env_a, task_a = make_env("Cartpole", num_envs=n, seed=9)
    env_b, task_b = make_env("Cartpole", num_envs=n, seed=9)
    env_a.reset(); env_b.reset()
    env_b.enable_cuda_graph()
    assert env_b.graph_kernels == 3
    g = torch.Generator().manual_seed(0)
    resets = 0
    for step in range(60):
        a = (torch.rand(n, 1, generator=g) * 2.4 - 1.2).to(task_a.device)
        oa, ra, da, _ = env_a.step(a)
        ob, rb, db, _ = env_b.step(a)
        assert torch.equal(oa["obs"], ob["obs"]), f"obs differ at step {step}"
        assert torch.equal(ra, rb) and torch.equal(da, db)
        resets += int(da.sum())
    assert resets > n // 8
    sa, sb = _snap(task_a), _snap(task_b)
    for k in sa:
        assert torch.equal(sa[k], sb[k]), k


def test_physics_plausibility_and_energy(cuda_lib):
    """Unactuated poles fall: from |theta0| <= pi/8 nearly every pole passes pi/2 within 300 steps and the env resets."""
    n = 1024
    env, task = make_env("Cartpole", num_envs=n, seed=0)
    env.reset()
    done_any = torch.zeros(n, dtype=torch.bool, device=task.device)
    for _ in range(300):
        _, _, d, _ = env.step(torch.zeros(n, 1, device=task.device))
        done_any |= d.bool()
    assert float(done_any.float().mean()) > 0.95
    # the push acts on the pole with the right sign through the prismatic joint: a pole leaning toward -y
    # (theta > 0, hinge about +x) is caught by accelerating the cart toward -y
    finals = []
    for push in (0.0, -0.05):
        env, task = make_env("Cartpole", num_envs=64, seed=0)
        env.reset()
        v = task.get_robot()
        v.dof_pos[:, 0], v.dof_pos[:, 1] = 0.0, 0.1
        v.dof_vel.zero_()
        for _ in range(10):
            env.step(torch.full((64, 1), push, device=task.device))
        assert int(task.reset_buf.sum()) == 0
        finals.append(v.dof_pos.clone())
    assert float(finals[0][:, 1].min()) > 0.1                        # unactuated: keeps falling
    assert float(finals[1][:, 1].max()) < float(finals[0][:, 1].min())  # pushed: falls less / recovers
    assert float(finals[1][:, 0].max()) < -0.01                      # the cart moved toward -y
