"""GPU: K1 (fused obs / reward / termination kernel) through the C ABI against the oracle.

Bar (BASELINE.json north_star): floats within 1e-5 relative error in fp32 (tolerances in
helpers.py), reset indices bit-exact.  Sizes the oracle finishes in seconds are compared
directly; BASELINE-scale sizes are checked through size-independent properties.
"""
import ctypes as C

import pytest
import torch

import helpers
from omniisaacgymenvs_b200 import _lib
from oracle import locomotion_oracle as lo

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("closed_form", [1, 0])
@pytest.mark.parametrize("name", helpers.golden_names())
def test_k1_matches_golden_reference_vectors(cuda_lib, name, closed_form):
    """Both angle-normalisation modes: 0 = literal atan2(sin, cos) chain, 1 = closed-form wrap (default)."""
    s, ref = helpers.load_golden(name)
    spec = lo.SPECS["Humanoid" if name.startswith("humanoid") else "Ant"]
    got = helpers.run_k1(spec, s, DEV, closed_form_angles=closed_form)
    helpers.assert_k1_matches(got, ref, name)


@pytest.mark.parametrize("robot", ["Humanoid", "Ant"])
def test_k1_closed_form_angles_agree_with_literal_chain(cuda_lib, robot):
    """The closed-form wrap must sit within a few ulp of the literal atan2(sin x, cos x) chain."""
    import math

    spec = lo.SPECS[robot]
    s = lo.synthetic_state(spec, 65536, 17)
    a = helpers.run_k1(spec, s, DEV, closed_form_angles=1)
    b = helpers.run_k1(spec, s, DEV, closed_form_angles=0)
    for k in a:
        if k != "obs":
            assert torch.equal(a[k], b[k]), k
    other = [c for c in range(spec.num_obs) if c not in (7, 8, 9)]
    assert torch.equal(a["obs"][:, other], b["obs"][:, other])
    d = (a["obs"][:, 7:10] - b["obs"][:, 7:10]).abs()
    d = torch.minimum(d, (2 * math.pi - d).abs())
    assert float(d.max()) <= 1.5e-6, float(d.max())


@pytest.mark.parametrize("robot", ["Humanoid", "Ant"])
@pytest.mark.parametrize("n", [1, 3, 31, 32, 33, 64, 127, 4096, 4099])
@pytest.mark.parametrize("tile", [0, 32, 64])
def test_k1_matches_oracle_all_tiles_and_ragged_sizes(cuda_lib, robot, n, tile):
    spec = lo.SPECS[robot]
    s = lo.synthetic_state(spec, n, seed=n + 11)
    ref = lo.post_physics_step(spec, {k: v.clone() for k, v in s.items()})
    got = helpers.run_k1(spec, s, DEV, tile=tile)
    helpers.assert_k1_matches(got, ref, f"{robot} n={n} tile={tile}")


@pytest.mark.parametrize("robot", ["Humanoid", "Ant"])
def test_k1_seeds_at_default_env_count(cuda_lib, robot):
    spec = lo.SPECS[robot]
    for seed in range(4):
        s = lo.synthetic_state(spec, 4096, seed)
        ref = lo.post_physics_step(spec, {k: v.clone() for k, v in s.items()})
        helpers.assert_k1_matches(helpers.run_k1(spec, s, DEV), ref, f"{robot} seed {seed}")


def test_k1_empty_input_is_a_noop(cuda_lib):
    spec = lo.HUMANOID
    s = lo.synthetic_state(spec, 4, 0)
    p = helpers.params_from_spec(spec, s["dof_lo"], s["dof_hi"])
    case = helpers.DeviceCase(spec, s, DEV)
    st, bf = case.structs()
    assert cuda_lib.oige_obs_reward_reset_f32(C.byref(p), C.byref(st), C.byref(bf), 0, None) == 0
    assert torch.isnan(case.outputs()["obs"]).all()


def test_k1_bad_arguments_return_errors_not_crashes(cuda_lib):
    spec = lo.HUMANOID
    s = lo.synthetic_state(spec, 8, 0)
    case = helpers.DeviceCase(spec, s, DEV)
    st, bf = case.structs()
    p = helpers.params_from_spec(spec, s["dof_lo"], s["dof_hi"])
    assert cuda_lib.oige_obs_reward_reset_f32(None, C.byref(st), C.byref(bf), 8, None) == -1
    p.num_obs = 86
    assert cuda_lib.oige_obs_reward_reset_f32(C.byref(p), C.byref(st), C.byref(bf), 8, None) == -2
    assert b"num_obs" in cuda_lib.oige_last_error()
    p.num_obs, p.num_dof = 87, 20
    assert cuda_lib.oige_obs_reward_reset_f32(C.byref(p), C.byref(st), C.byref(bf), 8, None) == -2
    p = helpers.params_from_spec(spec, s["dof_lo"], s["dof_hi"])
    bf.obs = None
    assert cuda_lib.oige_obs_reward_reset_f32(C.byref(p), C.byref(st), C.byref(bf), 8, None) == -1


@pytest.mark.parametrize("robot", ["Humanoid", "Ant"])
def test_k1_unaligned_views_take_the_plain_path(cuda_lib, robot):
    """Rows [1:] of every tensor: base pointers are no longer 16-byte aligned."""
    spec = lo.SPECS[robot]
    n = 257
    s = lo.synthetic_state(spec, n + 1, 5)
    sub = {k: (v[1:].clone() if v.shape[:1] == (n + 1,) else v) for k, v in s.items()}
    ref = lo.post_physics_step(spec, {k: v.clone() for k, v in sub.items()})
    dev = torch.device(DEV)
    big = helpers.DeviceCase(spec, s, DEV)
    st = _lib.OigeState(big.pos[1:].data_ptr(), big.quat[1:].data_ptr(), big.vel[1:].data_ptr(),
                        big.dof_pos[1:].data_ptr(), big.dof_vel[1:].data_ptr(), big.sensors[1:].data_ptr(), None)
    bf = _lib.OigeTaskBuffers(big.obs[1:].data_ptr(), big.rew[1:].data_ptr(), big.reset_buf[1:].data_ptr(),
                              big.progress_buf[1:].data_ptr(), big.potentials[1:].data_ptr(),
                              big.prev_potentials[1:].data_ptr(), big.up_vec[1:].data_ptr(),
                              big.heading_vec[1:].data_ptr(), big.actions[1:].data_ptr(), None, None)
    p = helpers.params_from_spec(spec, s["dof_lo"], s["dof_hi"])
    _lib.check(cuda_lib.oige_obs_reward_reset_f32(C.byref(p), C.byref(st), C.byref(bf), n,
                                                  _lib.current_stream_ptr(dev)), "k1")
    out = big.outputs()
    got = {k: v[1:] for k, v in out.items()}
    helpers.assert_k1_matches(got, ref, "unaligned")
    assert torch.isnan(out["obs"][0]).all(), "row 0 must be untouched"


def test_k1_clip_obs_clamps_outputs_not_metrics(cuda_lib):
    spec = lo.HUMANOID
    s = lo.synthetic_state(spec, 512, 3)
    ref = lo.post_physics_step(spec, {k: v.clone() for k, v in s.items()})
    got = helpers.run_k1(spec, s, DEV, clip_obs=0.5)
    assert float(got["obs"].abs().max()) <= 0.5
    torch.testing.assert_close(got["obs"], ref["obs"].clamp(-0.5, 0.5), rtol=1e-5, atol=1e-5)
    # reward / reset are computed from the unclamped values, like the reference (the clamp of
    # envs/vec_env_rlgames.py:42 acts on a copy after calculate_metrics)
    assert torch.equal(got["reset_buf"], ref["reset_buf"])
    torch.testing.assert_close(got["rew"], ref["rew"], rtol=helpers.REW_RTOL, atol=helpers.REW_ATOL)


def test_k1_progress_increment_flag(cuda_lib):
    spec = lo.ANT
    s = lo.synthetic_state(spec, 100, 2)
    got = helpers.run_k1(spec, s, DEV, progress_increment=0)
    assert torch.equal(got["progress_buf"], s["progress_buf"])
    got = helpers.run_k1(spec, s, DEV, progress_increment=1)
    assert torch.equal(got["progress_buf"], s["progress_buf"] + 1)


def test_k1_episode_stats_match_outputs(cuda_lib):
    spec = lo.HUMANOID
    s = lo.synthetic_state(spec, 5000, 9)
    got = helpers.run_k1(spec, s, DEV, with_stats=True)
    st = got["stats"]
    assert abs(float(st[0]) - float(got["rew"].double().sum())) < 1e-2
    assert int(st[1]) == int(got["reset_buf"].sum())
    assert int(st[2]) == int(got["progress_buf"][got["reset_buf"] != 0].sum())
    assert int(st[3]) == 5000


@pytest.mark.parametrize("robot,n", [("Humanoid", 16384), ("Humanoid", 65536), ("Ant", 65536), ("Humanoid", 1 << 20)])
def test_k1_full_size_properties(cuda_lib, robot, n):
    """BASELINE.json sizes (16k, 64k, 1M envs): size-independent properties instead of the oracle.
    (1) the result of a big launch equals the results of its 4096-row chunks launched separately,
    bit for bit (tile / grid independence); (2) a 4096-row sample matches the oracle; (3) reset
    flags equal the closed-form predicate on the inputs; (4) obs[:,0] is a copy of the height."""
    spec = lo.SPECS[robot]
    s = lo.synthetic_state(spec, n, 21)
    got = helpers.run_k1(spec, s, DEV)
    # (1)+(2) on three chunks spread over the range
    for start in (0, (n // 2 // 4096) * 4096, n - 4096):
        sub = {k: (v[start:start + 4096].clone() if v.shape[:1] == (n,) else v) for k, v in s.items()}
        part = helpers.run_k1(spec, sub, DEV, tile=32)
        for k in part:
            assert torch.equal(part[k], got[k][start:start + 4096]), f"{k} differs between big and chunked launch"
    sub = {k: (v[:4096].clone() if v.shape[:1] == (n,) else v) for k, v in s.items()}
    ref = lo.post_physics_step(spec, sub)
    helpers.assert_k1_matches({k: v[:4096] for k, v in got.items()}, ref, "sample")
    # (3)
    fallen = s["pos"][:, 2] < spec.termination_height
    timeout = (s["progress_buf"] + 1) >= spec.max_episode_length - 1
    assert torch.equal(got["reset_buf"], (fallen | timeout | (s["reset_buf"] != 0)).long())
    assert torch.equal(got["rew"][fallen], torch.full((int(fallen.sum()),), spec.death_cost))
    # (4)
    assert torch.equal(got["obs"][:, 0], s["pos"][:, 2])
    assert torch.isfinite(got["obs"]).all() and torch.isfinite(got["rew"]).all()


def test_k1_is_idempotent_on_pure_outputs(cuda_lib):
    """Running the pass twice on the same state (progress_increment=0) rewrites identical obs."""
    spec = lo.HUMANOID
    s = lo.synthetic_state(spec, 2048, 4)
    a = helpers.run_k1(spec, s, DEV, progress_increment=0)
    s2 = dict(s)
    s2["potentials"] = a["potentials"]
    s2["reset_buf"] = a["reset_buf"]
    b = helpers.run_k1(spec, s2, DEV, progress_increment=0)
    assert torch.equal(a["obs"], b["obs"])
    assert torch.equal(a["potentials"], b["potentials"])
    assert torch.equal(b["prev_potentials"], a["potentials"])


@pytest.mark.parametrize("robot", ["Humanoid", "Ant"])
def test_k1_fast_division_is_correctly_rounded(cuda_lib, robot):
    """unscale() uses a reciprocal + one-FMA-correction division (csrc: div_rn); it must equal the
    IEEE quotient torch computes, bit for bit, because the Ant limit cost thresholds it at 0.99."""
    spec = lo.SPECS[robot]
    n = 1 << 18
    s = lo.synthetic_state(spec, n, 33)
    got = helpers.run_k1(spec, s, DEV)
    d = spec.num_dof
    ref = (2.0 * s["dof_pos"] - s["dof_hi"] - s["dof_lo"]) / (s["dof_hi"] - s["dof_lo"])
    assert torch.equal(got["obs"][:, 12:12 + d], ref)
    assert torch.equal(got["obs"][:, 12 + d:12 + 2 * d], s["dof_vel"] * spec.dof_vel_scale)
    assert torch.equal(got["obs"][:, 12 + 2 * d + 6 * spec.num_sensors:], s["actions"])
    assert torch.equal(got["obs"][:, 12 + 2 * d:12 + 2 * d + 6 * spec.num_sensors],
                       s["sensors"].reshape(n, -1) * spec.contact_force_scale)


def test_k1_measured_deviation_table_is_inside_the_bar(cuda_lib):
    """The table DESIGN.md section 4 quotes (tests/k1_deviation.py; 4096 envs x 4 seeds per robot + the goldens):
    every output except the three angle columns and the reward is BIT-identical to the oracle (itself bit-identical
    to the verbatim reference functions); angles within 1.5e-6 on the circle; reward within 1e-6 relative."""
    import k1_deviation as kd

    for cases in (kd.synthetic_cases(), kd.golden_cases()):
        for robot, rows in kd.measure(cases).items():
            for name, a in rows.items():
                if "circle" in name:
                    assert a.max_abs <= helpers.ANGLE_TOL, (robot, name, a.max_abs)
                elif name.startswith("rew"):
                    assert a.max_rel <= 1e-6, (robot, name, a.max_rel)  # relative to max(|rew|, 1)
                else:
                    assert a.nbits == 0, (robot, name, a.nbits, a.max_abs)
