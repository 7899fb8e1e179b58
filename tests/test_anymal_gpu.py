"""GPU: the Anymal passes through the C ABI against oracle/anymal_oracle.py (pinned exactly to the verbatim reference
AnymalTask, tests/test_oracle_anymal.py) and the task / env classes end to end.

Bar: integer outputs (reset_buf, progress_buf) and everything that is a copy or a single rounded op (reset values,
commands, targets, last_actions, last_dof_vel, the dof / action / command observation columns) BIT-exact; observation
columns and rewards that go through quat_rotate / exp / a 12-term sum within 1e-6 (different summation order and
libm than torch's CPU kernels).  Physics (K2 with the PD drive) is covered against the float64 oracle below and in
tests/test_k2_dynamics_gpu.py."""
import ctypes as C
import os

import numpy as np
import pytest
import torch

import dyn_helpers as dh
import helpers
from omniisaacgymenvs_b200 import _lib
from omniisaacgymenvs_b200.robots.model import build_model
from omniisaacgymenvs_b200.utils.task_util import make_env
from oracle import anymal_oracle as ao
from oracle import philox

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL = dict(rtol=1e-6, atol=1e-6)


def make_params(model, inc=1, clip=5.0, spec=ao.ANYMAL):
    p = _lib.OigeAnymalParams()
    p.progress_increment = inc
    p.lin_vel_scale, p.ang_vel_scale, p.dof_pos_scale, p.dof_vel_scale = (spec.lin_vel_scale, spec.ang_vel_scale,
                                                                          spec.dof_pos_scale, spec.dof_vel_scale)
    p.action_scale, p.dt = spec.action_scale, spec.dt
    rs = spec.rew_scales()
    p.rew_lin_vel_xy, p.rew_ang_vel_z, p.rew_lin_vel_z = rs["lin_vel_xy"], rs["ang_vel_z"], rs["lin_vel_z"]
    p.rew_joint_acc, p.rew_action_rate, p.rew_cosmetic = rs["joint_acc"], rs["action_rate"], rs["cosmetic"]
    for k in range(2):
        p.command_x[k], p.command_y[k], p.command_yaw[k] = (spec.command_x_range[k], spec.command_y_range[k],
                                                            spec.command_yaw_range[k])
    p.fallen_threshold, p.max_episode_length, p.clip_obs = spec.fallen_threshold, float(spec.max_episode_length), clip
    lim = model.dof_limits()
    d = spec.default_dof_pos()
    for k in range(12):
        p.default_dof_pos[k], p.dof_limits_lower[k], p.dof_limits_upper[k] = float(d[k]), float(lim[k, 0]), float(lim[k, 1])
    return p


class Case:
    """Device copies of an oracle state dict + the ABI structs."""

    KEYS = ("pos", "quat", "vel", "dof_pos", "dof_vel", "actions", "last_actions", "last_dof_vel", "commands")

    def __init__(self, s, n):
        self.n = n
        z = lambda *shape, dt=torch.float32: torch.zeros(*shape, dtype=dt, device=DEV)
        for k in self.KEYS:
            setattr(self, k, s[k].to(DEV).contiguous() if k in s else z(n, {"pos": 3, "quat": 4, "vel": 6, "commands": 3}.get(k, 12)))
        self.targets = s["current_targets"].to(DEV).contiguous() if "current_targets" in s else z(n, 12)
        self.obs, self.rew = z(n, 48), z(n)
        self.reset = s["reset_buf"].to(DEV).contiguous() if "reset_buf" in s else z(n, dt=torch.int64)
        self.progress = s["progress_buf"].to(DEV).contiguous() if "progress_buf" in s else z(n, dt=torch.int64)
        self.stats = z(_lib.STATS_SLOTS, 4, dt=torch.float64)
        self.init_pos = s["initial_root_pos"].to(DEV).contiguous() if "initial_root_pos" in s else self.pos.clone()
        self.init_rot = s["initial_root_rot"].to(DEV).contiguous() if "initial_root_rot" in s else self.quat.clone()
        self.state = _lib.OigeState(self.pos.data_ptr(), self.quat.data_ptr(), self.vel.data_ptr(), self.dof_pos.data_ptr(),
                                    self.dof_vel.data_ptr(), None, self.targets.data_ptr())
        self.buf = _lib.OigeAnymalBuffers(self.obs.data_ptr(), self.rew.data_ptr(), self.reset.data_ptr(),
                                          self.progress.data_ptr(), self.actions.data_ptr(), self.last_actions.data_ptr(),
                                          self.last_dof_vel.data_ptr(), self.commands.data_ptr(), self.targets.data_ptr(),
                                          self.stats.data_ptr(), None)
        self.init = _lib.OigeResetInit(self.init_pos.data_ptr(), self.init_rot.data_ptr(), None)

    def post(self, lib, p):
        _lib.check(lib.oige_anymal_obs_reward_reset_f32(C.byref(p), C.byref(self.state), C.byref(self.buf), self.n,
                                                        _lib.current_stream_ptr(DEV)), "post")
        torch.cuda.synchronize()


def assert_post_matches(c, out, clip=5.0):
    obs = torch.clamp(out["obs"], -clip, clip)
    got = c.obs.cpu()
    assert torch.equal(got[:, 9:], obs[:, 9:]), "commands / dof / action columns are single-op: bit-exact"
    torch.testing.assert_close(got[:, :9], obs[:, :9], **TOL)
    torch.testing.assert_close(c.rew.cpu(), out["rew"], **TOL)
    assert torch.equal(c.rew.cpu() == -1.0, out["rew"] == -1.0) and float(c.rew.min()) >= -1.0
    assert torch.equal(c.reset.cpu(), out["reset_buf"]) and torch.equal(c.progress.cpu(), out["progress_buf"])
    assert torch.equal(c.last_actions.cpu(), out["last_actions"]) and torch.equal(c.last_dof_vel.cpu(), out["last_dof_vel"])


@pytest.mark.parametrize("n", [1, 33, 4096, 100001])
def test_obs_reward_done_match_the_oracle(cuda_lib, n):
    model = build_model("Anymal")
    s = ao.synthetic_state(ao.ANYMAL, n, 3)
    s["progress_buf"][::7] = 2998   # -> 2999 after the increment: the episode limit
    s["progress_buf"][1::7] = 2997
    s["reset_buf"] = torch.ones(n, dtype=torch.int64)  # incoming flags are NOT carried over (anymal.py:293)
    c = Case(s, n)
    c.post(cuda_lib, make_params(model))
    out = ao.post_physics_step(ao.ANYMAL, {k: v.clone() for k, v in s.items()})
    assert_post_matches(c, out)
    st = c.stats.sum(0).cpu()
    assert int(st[3]) == n and int(st[1]) == int(out["reset_buf"].sum())
    assert abs(float(st[0]) - float(out["rew"].double().sum())) < 1e-5 * max(1.0, n)


def test_unaligned_views_take_the_plain_path(cuda_lib):
    """Tensors offset by one row are not 16-byte aligned for every field: the cooperative path must agree."""
    model = build_model("Anymal")
    n = 257
    s = ao.synthetic_state(ao.ANYMAL, n + 1, 4)
    c_full = Case(s, n + 1)
    sub = {k: (v[1:] if torch.is_tensor(v) and v.shape[0] == n + 1 else v) for k, v in s.items()}
    c = Case.__new__(Case)
    c.n = n
    for k in Case.KEYS + ("targets", "obs", "rew", "reset", "progress", "init_pos", "init_rot"):
        setattr(c, k, getattr(c_full, k)[1:])
    c.stats = c_full.stats
    c.state = _lib.OigeState(c.pos.data_ptr(), c.quat.data_ptr(), c.vel.data_ptr(), c.dof_pos.data_ptr(),
                             c.dof_vel.data_ptr(), None, c.targets.data_ptr())
    c.buf = _lib.OigeAnymalBuffers(c.obs.data_ptr(), c.rew.data_ptr(), c.reset.data_ptr(), c.progress.data_ptr(),
                                   c.actions.data_ptr(), c.last_actions.data_ptr(), c.last_dof_vel.data_ptr(),
                                   c.commands.data_ptr(), c.targets.data_ptr(), None, None)
    c.post(cuda_lib, make_params(model))
    out = ao.post_physics_step(ao.ANYMAL, {k: v.clone() for k, v in sub.items()})
    assert_post_matches(c, out)
    assert float(c_full.obs[0].abs().sum()) == 0.0  # row 0 untouched


def test_reference_episode_through_the_c_abi(cuda_lib):
    """tests/golden/episode_anymal.npz (recorded from the verbatim reference classes) replayed kernel by kernel."""
    z = np.load(os.path.join(helpers.GOLDEN_DIR, "episode_anymal.npz"))
    ep = {k: torch.from_numpy(z[k]) for k in z.files}
    steps, n = ep["actions"].shape[:2]
    model = build_model("Anymal")
    assert torch.equal(torch.from_numpy(model.dof_limits()[:, 0]), ep["dof_lo"])
    lib, p, s = cuda_lib, make_params(model), _lib.current_stream_ptr(DEV)
    c = Case({"pos": ep["spawn_pos"].clone(), "quat": ep["initial_root_rot"].clone(), "reset_buf": torch.ones(n, dtype=torch.int64),
              "initial_root_pos": ep["initial_root_pos"], "initial_root_rot": ep["initial_root_rot"],
              "current_targets": ep["default_dof_pos"].clone()}, n)
    ids = torch.arange(n, device=DEV)
    rv, rc_ = ep["post_reset_rand_vel"].to(DEV).contiguous(), ep["post_reset_rand_cmd"].to(DEV).contiguous()
    _lib.check(lib.oige_anymal_reset_idx_f32(C.byref(p), C.byref(c.state), C.byref(c.buf), C.byref(c.init), ids.data_ptr(), n,
                                             rv.data_ptr(), rc_.data_ptr(), 0, 0, 0, s), "reset_idx")
    assert torch.equal(c.dof_pos.cpu(), ep["post_reset_dof_pos"]) and torch.equal(c.dof_vel.cpu(), ep["post_reset_dof_vel"])
    assert torch.equal(c.commands.cpu(), ep["post_reset_commands"]) and int(c.reset.sum()) == 0
    for t in range(steps):
        assert torch.equal(c.reset.cpu() != 0, ep["reset_mask"][t]), f"step {t}: reset set"
        a = ep["actions"][t].to(DEV).contiguous()
        rv, rc_ = ep["rand_vel"][t].to(DEV).contiguous(), ep["rand_cmd"][t].to(DEV).contiguous()
        _lib.check(lib.oige_anymal_reset_masked_f32(C.byref(p), C.byref(c.state), C.byref(c.buf), C.byref(c.init),
                                                    a.data_ptr(), 1.0, rv.data_ptr(), rc_.data_ptr(), 0, 0, 0, n, s), "pre")
        for got, key in ((c.dof_pos, "after_reset_dof_pos"), (c.dof_vel, "after_reset_dof_vel"), (c.pos, "after_reset_pos"),
                         (c.quat, "after_reset_quat"), (c.vel, "after_reset_vel"), (c.commands, "commands"),
                         (c.targets, "targets"), (c.last_actions, "after_reset_last_actions"),
                         (c.last_dof_vel, "after_reset_last_dof_vel"), (c.actions, "actions")):
            assert torch.equal(got.cpu(), ep[key][t]), f"step {t}: {key}"
        assert int(c.reset.sum()) == 0
        for dst, key in ((c.pos, "inj_pos"), (c.quat, "inj_quat"), (c.vel, "inj_vel"), (c.dof_pos, "inj_dof_pos"),
                         (c.dof_vel, "inj_dof_vel")):
            dst.copy_(ep[key][t])
        if t == 2:
            c.progress[: n // 4] = ao.ANYMAL.max_episode_length - 3
        c.post(lib, p)
        assert_post_matches(c, {k: ep[k][t] for k in ("obs", "rew", "reset_buf", "progress_buf", "last_actions", "last_dof_vel")})


def test_philox_reset_is_bit_exact_and_sharding_invariant(cuda_lib):
    n, seed, counter, off = 1000, 0xFEED5EED1234, 9, 70000
    model = build_model("Anymal")
    s = ao.synthetic_state(ao.ANYMAL, n, 8)
    s["reset_buf"] = (torch.arange(n) % 3 != 0).long()
    s["current_targets"] = s["dof_pos"].clone()
    s["initial_root_pos"] = torch.randn(n, 3)
    s["initial_root_rot"] = torch.tensor([[1.0, 0, 0, 0]]).repeat(n, 1)
    before = {k: v.clone() for k, v in s.items()}
    c = Case(s, n)
    a = (torch.rand(n, 12) * 3 - 1.5).to(DEV)
    p = make_params(model)
    _lib.check(cuda_lib.oige_anymal_reset_masked_f32(C.byref(p), C.byref(c.state), C.byref(c.buf), C.byref(c.init), a.data_ptr(),
                                                     1.0, None, None, seed, counter, off, n, _lib.current_stream_ptr(DEV)), "pre")
    gids = np.arange(off, off + n)
    uv = torch.from_numpy(philox.reset_uniforms(seed, counter, gids, 12, 1))
    uc = torch.from_numpy(philox.reset_uniforms(seed, counter, gids, 3, 0))
    ref = {k: v.clone() for k, v in before.items()}
    ref["dof_lo"], ref["dof_hi"] = (torch.from_numpy(model.dof_limits()[:, k]) for k in (0, 1))
    ids = ref["reset_buf"].nonzero(as_tuple=False).squeeze(-1)
    ao.reset_idx(ao.ANYMAL, ref, ids, uv[ids], uc[ids])
    ao.integrate_targets(ao.ANYMAL, ref, torch.clamp(a.cpu(), -1.0, 1.0))
    for got, key in ((c.dof_pos, "dof_pos"), (c.dof_vel, "dof_vel"), (c.pos, "pos"), (c.quat, "quat"), (c.vel, "vel"),
                     (c.commands, "commands"), (c.targets, "current_targets"), (c.last_actions, "last_actions"),
                     (c.last_dof_vel, "last_dof_vel"), (c.actions, "actions"), (c.reset, "reset_buf"),
                     (c.progress, "progress_buf")):
        assert torch.equal(got.cpu(), ref[key]), key
    # a shard owning global envs [off+400, off+700) draws the same numbers
    s2 = {k: (v[400:700].clone() if torch.is_tensor(v) and v.shape[0] == n else v) for k, v in before.items()}
    c2 = Case(s2, 300)
    _lib.check(cuda_lib.oige_anymal_reset_masked_f32(C.byref(p), C.byref(c2.state), C.byref(c2.buf), C.byref(c2.init), None, 1.0,
                                                     None, None, seed, counter, off + 400, 300, _lib.current_stream_ptr(DEV)), "pre")
    assert torch.equal(c2.dof_vel.cpu(), ref["dof_vel"][400:700]) and torch.equal(c2.commands.cpu(), ref["commands"][400:700])
    # command ranges (Anymal.yaml:21-24)
    cm = ref["commands"][ids]
    assert float(cm[:, 0].abs().max()) <= 2.0 and float(cm[:, 1].abs().max()) <= 1.0 and float(cm[:, 0].std()) > 0.9


def test_argument_errors_come_back_as_codes(cuda_lib):
    model = build_model("Anymal")
    c = Case(ao.synthetic_state(ao.ANYMAL, 8, 0), 8)
    p = make_params(model)
    assert cuda_lib.oige_anymal_obs_reward_reset_f32(None, C.byref(c.state), C.byref(c.buf), 8, None) == -1
    bad = _lib.OigeAnymalBuffers(c.obs.data_ptr(), c.rew.data_ptr(), c.reset.data_ptr(), c.progress.data_ptr(), None, None,
                                 None, None, None, None, None)
    assert cuda_lib.oige_anymal_obs_reward_reset_f32(C.byref(p), C.byref(c.state), C.byref(bad), 8, None) == -1
    assert b"null task buffer" in cuda_lib.oige_last_error()
    st = _lib.OigeState(c.pos.data_ptr(), c.quat.data_ptr(), c.vel.data_ptr(), c.dof_pos.data_ptr(), c.dof_vel.data_ptr(), None, None)
    a = torch.zeros(8, 12, device=DEV)
    assert cuda_lib.oige_anymal_env_step_f32(1, C.byref(p), C.byref(st), C.byref(c.buf), C.byref(c.init), a.data_ptr(), 1.0, 0, 0, 0,
                                             8, 2, 0.01, None) == -1
    assert b"dof_pos_targets" in cuda_lib.oige_last_error()


# ---- task / env level ---------------------------------------------------------------------------------------
def _snap(task):
    v = task.get_robot()
    return {k: t.clone() for k, t in (("pos", v.root_pos), ("quat", v.root_quat), ("vel", v.root_vel), ("q", v.dof_pos),
                                      ("qd", v.dof_vel), ("obs", task.obs_buf), ("rew", task.rew_buf), ("reset", task.reset_buf),
                                      ("progress", task.progress_buf), ("targets", task.current_targets),
                                      ("commands", task.commands), ("last_actions", task.last_actions),
                                      ("last_dof_vel", task.last_dof_vel))}


def test_env_surface_matches_the_reference_task(cuda_lib):
    env, task = make_env("Anymal", num_envs=256, seed=0)
    assert task.num_observations == 48 and task.num_actions == 12 and task.max_episode_length == 3000
    assert task.clip_obs == 5.0 and task.control_frequency_inv == 2 and env._world.get_physics_dt() == 0.01
    v = task.get_robot()
    assert v.model.dof_names == ao.DOF_NAMES and float(v.root_pos[:, 2].min()) == pytest.approx(0.62)
    assert torch.equal(task.default_dof_pos[0].cpu(), ao.ANYMAL.default_dof_pos())
    assert task.current_targets.data_ptr() == v.dof_pos_targets.data_ptr()
    assert torch.equal(v.dof_pos, task.default_dof_pos) and float(v.dof_vel.abs().max()) <= 0.1
    assert float(task.commands[:, 0].abs().max()) <= 2.0 and float(task.commands[:, 1:].abs().max()) <= 1.0
    obs = env.reset()
    assert obs["obs"].shape == (256, 48) and float(obs["obs"].abs().max()) <= 5.0
    o, r, d, extras = env.step(torch.zeros(256, 12, device=task.device))
    assert o["obs"].shape == (256, 48) and r.shape == (256,) and d.dtype == torch.int64 and extras == {}
    assert not bool(v.is_base_below_threshold(0.51, 0.0).any())


def test_fused_step_equals_hook_by_hook_path(cuda_lib):
    from omniisaacgymenvs_b200.tasks.anymal import AnymalTask

    class Hooked(AnymalTask):  # overriding a hook turns the fused path off
        def is_done(self):
            AnymalTask.is_done(self)

    n = 600
    env_a, task_a = make_env("Anymal", num_envs=n, seed=3)
    env_b, task_b = make_env("Anymal", num_envs=n, seed=3)
    task_b.__class__ = Hooked
    assert task_a.fused_step_enabled() and not task_b.fused_step_enabled()
    env_a.reset(); env_b.reset()
    g = torch.Generator().manual_seed(0)
    resets = 0
    for step in range(150):
        a = (torch.rand(n, 12, generator=g) * 2.4 - 1.2).to(task_a.device)
        oa, ra, da, _ = env_a.step(a)
        ob, rb, db, _ = env_b.step(a)
        assert torch.equal(oa["obs"], ob["obs"]), f"step {step}"
        assert torch.equal(ra, rb) and torch.equal(da, db)
        resets += int(da.sum())
    assert resets > n // 4  # random targets make them fall
    sa, sb = _snap(task_a), _snap(task_b)
    for k in sa:
        assert torch.equal(sa[k], sb[k]), k
    assert torch.equal(task_b.fallen_over, task_b.get_robot().root_pos[:, 2] < 0.51)


def test_post_physics_step_matches_oracle_on_live_state(cuda_lib):
    n = 1024
    env, task = make_env("Anymal", num_envs=n, seed=5)
    env.reset()
    v = task.get_robot()
    for step in range(60):
        a = torch.rand(n, 12, device=task.device) * 2 - 1
        flagged = task.reset_buf.bool()
        prog = torch.where(flagged, torch.zeros_like(task.progress_buf), task.progress_buf).cpu()
        la = torch.where(flagged[:, None], torch.zeros_like(task.last_actions), task.last_actions).cpu()
        lq = torch.where(flagged[:, None], torch.zeros_like(task.last_dof_vel), task.last_dof_vel).cpu()
        o, r, d, _ = env.step(a)
        s = {"pos": v.root_pos.cpu(), "quat": v.root_quat.cpu(), "vel": v.root_vel.cpu(), "dof_pos": v.dof_pos.cpu(),
             "dof_vel": v.dof_vel.cpu(), "default_dof_pos": task.default_dof_pos.cpu(), "commands": task.commands.cpu(),
             "actions": a.cpu(), "last_actions": la, "last_dof_vel": lq, "progress_buf": prog}
        out = ao.post_physics_step(ao.ANYMAL, s)
        torch.testing.assert_close(o["obs"].cpu(), torch.clamp(out["obs"], -5.0, 5.0), **TOL)
        torch.testing.assert_close(r.cpu(), out["rew"], **TOL)
        assert torch.equal(d.cpu(), out["reset_buf"]) and torch.equal(task.progress_buf.cpu(), out["progress_buf"])


def test_sharded_envs_equal_rows_of_the_unsharded_run(cuda_lib):
    n = 512
    env_full, task_full = make_env("Anymal", num_envs=n, seed=11)
    env_half, task_half = make_env("Anymal", num_envs=n // 2, seed=11, env_id_offset=n // 2, global_num_envs=n)
    assert torch.equal(task_full.initial_root_pos[n // 2:], task_half.initial_root_pos)
    env_full.reset(); env_half.reset()
    g = torch.Generator().manual_seed(2)
    for step in range(80):
        a = (torch.rand(n, 12, generator=g) * 2 - 1).to(task_full.device)
        of, rf, df, _ = env_full.step(a)
        oh, rh, dh_, _ = env_half.step(a[n // 2:].contiguous())
        assert torch.equal(of["obs"][n // 2:], oh["obs"]), f"step {step}"
        assert torch.equal(rf[n // 2:], rh) and torch.equal(df[n // 2:], dh_)
    assert int(task_full.read_episode_stats(reset=False)[3]) == n * 81


@pytest.mark.parametrize("n", [512, 16384])
def test_cuda_graph_step_equals_eager_step(cuda_lib, n):
    env_a, task_a = make_env("Anymal", num_envs=n, seed=9)
    env_b, task_b = make_env("Anymal", num_envs=n, seed=9)
    env_a.reset(); env_b.reset()
    env_b.enable_cuda_graph()
    assert env_b.graph_kernels == 3
    g = torch.Generator().manual_seed(0)
    resets = 0
    for step in range(80):
        a = (torch.rand(n, 12, generator=g) * 2.4 - 1.2).to(task_a.device)
        oa, ra, da, _ = env_a.step(a)
        ob, rb, db, _ = env_b.step(a)
        assert torch.equal(oa["obs"], ob["obs"]), f"obs differ at step {step}"
        assert torch.equal(ra, rb) and torch.equal(da, db)
        resets += int(da.sum())
    assert resets > n // 8
    sa, sb = _snap(task_a), _snap(task_b)
    for k in sa:
        assert torch.equal(sa[k], sb[k]), k


def test_pd_drive_holds_the_stance_and_tracks_targets_like_the_float64_oracle(cuda_lib):
    """Zero actions: the PD drive holds the default stance, nobody falls, the base settles well above the 0.51 m
    threshold.  The same 40 control steps (80 substeps of 0.01 s, foot contacts, saturating drives) in the float64
    oracle end within millimetres."""
    n = 256
    env, task = make_env("Anymal", num_envs=n, seed=0)
    env.reset()
    v = task.get_robot()
    st = {"pos": v.root_pos.cpu().numpy(), "quat": v.root_quat.cpu().numpy(), "vel": v.root_vel.cpu().numpy(),
          "q": v.dof_pos.cpu().numpy(), "qd": v.dof_vel.cpu().numpy()}
    for _ in range(200):
        _, _, d, _ = env.step(torch.zeros(n, 12, device=task.device))
        assert int(d.sum()) == 0
    z = v.root_pos[:, 2]
    assert 0.53 < float(z.min()) and float(z.max()) < 0.60
    assert float((v.dof_pos - task.default_dof_pos).abs().max()) < 0.02
    # oracle: same start (after the first pre-physics reset the state equals the post_reset one), 40 steps
    env2, task2 = make_env("Anymal", num_envs=n, seed=0)
    env2.reset()
    v2 = task2.get_robot()
    start = {"pos": v2.root_pos.cpu().numpy(), "quat": v2.root_quat.cpu().numpy(), "vel": v2.root_vel.cpu().numpy(),
             "q": v2.dof_pos.cpu().numpy(), "qd": v2.dof_vel.cpu().numpy()}
    from oracle import dynamics as dyn

    o = {k: np.ascontiguousarray(a.astype(np.float64)) for k, a in start.items()}
    tg = np.ascontiguousarray(task2.current_targets.cpu().numpy().astype(np.float64))
    desc = dyn.make_desc(v2.model)
    for _ in range(40):
        env2.step(torch.zeros(n, 12, device=task2.device))
        assert dyn.step(desc, o["pos"], o["quat"], o["vel"], o["q"], o["qd"], np.zeros((n, 12)), None, 2, 0.01, 8, targets=tg) == 0
    assert np.abs(v2.root_pos.cpu().numpy() - o["pos"]).max() < 5e-3
    assert np.abs(v2.dof_pos.cpu().numpy() - o["q"]).max() < 5e-3
