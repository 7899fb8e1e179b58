"""GPU: K2 (warp-per-env articulated-body step, fp32) through the C ABI against the float64 CPU
reference integrator (oracle/dynamics_oracle.c -- a different algorithm: dense projection method +
Cholesky).  PhysX itself is unavailable (documented substitution); tolerances per tier:

  single substep               |dq|, |dpos| <= 2e-5, |dqd| <= 2e-3 * max(1, |qd|) (fp32 rounding only)
  64 substeps, free flight     joint positions within 2e-3 rad, root within 2e-3 m
  64 substeps, ground contact  statistical: >= 97 % of envs within 2e-2 (contact switching is
                               chaotic; fp32 vs fp64 trajectories legitimately diverge in a few)
"""
import numpy as np
import pytest

import dyn_helpers as dh

pytestmark = pytest.mark.gpu
DT = 0.0083


def _set_mode(monkeypatch, mode):
    """The library reads its OIGE_* knobs once: set them, then ask it to re-read."""
    from omniisaacgymenvs_b200 import _lib

    if mode.startswith("role"):
        monkeypatch.setenv("OIGE_K2_MODE", "role")
        monkeypatch.setenv("OIGE_K2_EPC", "128" if mode == "role128" else "32")
    else:
        monkeypatch.setenv("OIGE_K2_MODE", mode)
        monkeypatch.delenv("OIGE_K2_EPC", raising=False)
    _lib.reload_env()


@pytest.fixture(autouse=True, params=["warp", "thread", "role", "role128"])
def k2_kernel(request, monkeypatch):
    """Every test runs against ALL articulation kernels: warp-per-env (lane = link, shuffles), thread-per-env
    (lane = env, local-memory link arrays) and the role-split whole-step kernel K2-F in physics-only mode, with 32-env
    and with 128-env CTAs.  OIGE_K2_MODE / OIGE_K2_EPC override the size-based choice."""
    from omniisaacgymenvs_b200 import _lib

    _set_mode(monkeypatch, request.param)
    yield request.param
    monkeypatch.undo()
    _lib.reload_env()


def _err(a, b):
    return np.abs(a - b).max()


def _quat_err(a, b):
    return np.minimum(np.abs(a - b).max(axis=1), np.abs(a + b).max(axis=1)).max()


@pytest.mark.parametrize("robot", ["Humanoid", "Ant"])
def test_single_substep_free_flight(cuda_lib, robot):
    model, spec = dh.specs(robot)
    st = dh.random_state(model, 512, 0, height=5.0)
    g = dh.run_gpu(cuda_lib, model, spec, st, 1, DT)
    o = dh.run_oracle(model, spec, st, 1, DT)
    assert _err(g["q"], o["q"]) < 2e-5
    assert _err(g["pos"], o["pos"]) < 2e-5
    assert _quat_err(g["quat"], o["quat"]) < 2e-5
    scale = np.maximum(1.0, np.abs(o["qd"]))
    assert (np.abs(g["qd"] - o["qd"]) / scale).max() < 2e-3
    assert (np.abs(g["vel"] - o["vel"]) / np.maximum(1.0, np.abs(o["vel"]))).max() < 2e-3


@pytest.mark.parametrize("robot", ["Humanoid", "Ant"])
def test_single_substep_with_ground_contact(cuda_lib, robot):
    model, spec = dh.specs(robot)
    # low enough that feet / legs penetrate the ground in most envs
    h = 1.25 if robot == "Humanoid" else 0.3
    st = dh.random_state(model, 512, 1, height=h, tilt=0.15)
    g = dh.run_gpu(cuda_lib, model, spec, st, 1, DT)
    o = dh.run_oracle(model, spec, st, 1, DT)
    assert np.abs(o["sensors"]).max() > 1.0, "test must actually exercise contacts"
    assert _err(g["q"], o["q"]) < 5e-5
    assert _err(g["pos"], o["pos"]) < 5e-5
    scale = np.maximum(1.0, np.abs(o["qd"]))
    assert (np.abs(g["qd"] - o["qd"]) / scale).max() < 5e-3
    fs = np.maximum(10.0, np.abs(o["sensors"]))
    assert (np.abs(g["sensors"] - o["sensors"]) / fs).max() < 5e-3


@pytest.mark.parametrize("robot", ["Humanoid", "Ant"])
def test_rollout_64_free_flight(cuda_lib, robot):
    # smooth dynamics only: limb-limb contacts switch chaotically like ground contacts do; their 32-substep rollout has
    # its own (statistical) test below
    model, spec = dh.specs(robot, enable_self_collisions=False)
    st = dh.random_state(model, 256, 2, height=50.0, qd_scale=0.5)
    g = dh.run_gpu(cuda_lib, model, spec, st, 2, DT, calls=32)
    o = dh.run_oracle(model, spec, st, 2, DT, calls=32)
    assert _err(g["q"], o["q"]) < 2e-3
    assert _err(g["pos"], o["pos"]) < 2e-3
    assert _quat_err(g["quat"], o["quat"]) < 2e-3


@pytest.mark.parametrize("robot", ["Humanoid", "Ant"])
def test_rollout_64_with_contacts_statistical(cuda_lib, robot):
    model, spec = dh.specs(robot)
    h = model.spawn_height
    st = dh.random_state(model, 1024, 3, height=h, qd_scale=0.2, root_vel_scale=0.1, tilt=0.05)
    g = dh.run_gpu(cuda_lib, model, spec, st, 2, DT, calls=32)
    o = dh.run_oracle(model, spec, st, 2, DT, calls=32)
    for k in g:
        assert np.isfinite(g[k]).all(), k
    per_env = np.maximum(np.abs(g["q"] - o["q"]).max(axis=1), np.abs(g["pos"] - o["pos"]).max(axis=1))
    frac_ok = float((per_env < 2e-2).mean())
    assert frac_ok >= 0.97, f"only {frac_ok:.3f} of envs within 2e-2 after 64 substeps (median {np.median(per_env):.2e})"
    assert np.median(per_env) < 2e-3


@pytest.mark.parametrize("robot", ["Humanoid", "Ant"])
def test_pd_position_drive_matches_oracle(cuda_lib, robot):
    """PD position-target actuation (set_joint_position_targets): implicit drive, saturating at max force.
    (A synthetic drive on the Humanoid tree: no self-collision -- K2-F has no position-target variant for 3 roles.)"""
    model, spec = dh.specs(robot, enable_self_collisions=False)
    model.drive_stiffness[1:] = 60.0
    model.drive_damping[1:] = 3.0
    model.drive_max_force[1:] = 40.0  # some joints saturate with the random targets below
    st = dh.random_state(model, 512, 21, height=5.0)
    rng = np.random.default_rng(3)
    lim = model.dof_limits()
    st["targets"] = np.ascontiguousarray((lim[:, 0] + rng.random((512, model.num_dof)) * (lim[:, 1] - lim[:, 0])).astype(np.float32))
    st["actions"][:] *= 0.2
    g = dh.run_gpu(cuda_lib, model, spec, st, 2, DT, calls=4)
    o = dh.run_oracle(model, spec, st, 2, DT, calls=4)
    assert _err(g["q"], o["q"]) < 1e-4 and _err(g["pos"], o["pos"]) < 1e-4
    assert (np.abs(g["qd"] - o["qd"]) / np.maximum(1.0, np.abs(o["qd"]))).max() < 5e-3
    # the drive does something: without targets the joints end elsewhere
    st2 = {k: v for k, v in st.items() if k != "targets"}
    g2 = dh.run_gpu(cuda_lib, model, spec, st2, 2, DT, calls=4)
    assert _err(g["q"], g2["q"]) > 1e-2


def test_both_kernels_agree_with_each_other(cuda_lib, monkeypatch):
    """warp-per-env vs thread-per-env on the same state: same model, different summation orders."""
    model, spec = dh.specs("Humanoid", enable_self_collisions=False)  # (only K2-F integrates self-collision)
    st = dh.random_state(model, 256, 11, height=1.3, tilt=0.1)
    _set_mode(monkeypatch, "warp")
    a = dh.run_gpu(cuda_lib, model, spec, st, 2, DT)
    _set_mode(monkeypatch, "thread")
    b = dh.run_gpu(cuda_lib, model, spec, st, 2, DT)
    assert _err(a["q"], b["q"]) < 5e-5 and _err(a["pos"], b["pos"]) < 5e-5
    assert (np.abs(a["qd"] - b["qd"]) / np.maximum(1.0, np.abs(a["qd"]))).max() < 5e-3
    for mode in ("role", "role128"):  # role split: trunk / limbs on separate warps, link data in shared memory
        _set_mode(monkeypatch, mode)
        c = dh.run_gpu(cuda_lib, model, spec, st, 2, DT)
        assert _err(c["q"], b["q"]) < 5e-5 and _err(c["pos"], b["pos"]) < 5e-5
        assert (np.abs(c["qd"] - b["qd"]) / np.maximum(1.0, np.abs(b["qd"]))).max() < 5e-3
        fs = np.maximum(10.0, np.abs(b["sensors"]))
        assert (np.abs(c["sensors"] - b["sensors"]) / fs).max() < 5e-3


@pytest.mark.parametrize("robot,n", [("Humanoid", 65536), ("Ant", 262144)])
def test_full_size_rows_are_independent(cuda_lib, robot, n, monkeypatch):
    """BASELINE.json sizes (65536 Humanoids, 262144 Ants), size-independent properties of the batched step: envs
    never interact, so (1) stepping a PERMUTATION of the envs gives the permuted result bit for bit, (2) a slice
    stepped alone equals the rows of the full launch, (3) a random sample of 512 rows agrees with the float64
    oracle after one substep, (4) everything stays finite over 16 substeps in contact."""
    # the thread-per-env kernel has a many-envs and a single-wave instantiation (different ptxas schedules): pin one,
    # as the library does for the chunks of one step
    monkeypatch.setenv("OIGE_K2_MINB", "4")
    from omniisaacgymenvs_b200 import _lib
    _lib.reload_env()
    model, spec = dh.specs(robot)
    st = dh.random_state(model, n, 21, height=model.spawn_height * 0.8, tilt=0.2)  # feet in / near the ground
    full = dh.run_gpu(cuda_lib, model, spec, st, 2, DT)
    rng = np.random.default_rng(0)
    perm = rng.permutation(n)
    shuffled = dh.run_gpu(cuda_lib, model, spec, {k: np.ascontiguousarray(v[perm]) for k, v in st.items()}, 2, DT)
    for k in full:
        assert np.array_equal(shuffled[k], full[k][perm]), k
    a, b = n // 2 - 1000, n // 2 + 1048
    part = dh.run_gpu(cuda_lib, model, spec, {k: np.ascontiguousarray(v[a:b]) for k, v in st.items()}, 2, DT)
    import os
    same_kernel = (os.environ.get("OIGE_K2_MODE") in ("warp", "thread", "role"))  # forced: slice and whole use one kernel
    for k in full:
        if same_kernel:
            assert np.array_equal(part[k], full[k][a:b]), k
    pick = np.sort(rng.choice(n, 512, replace=False))
    sub = {k: np.ascontiguousarray(v[pick]) for k, v in st.items()}
    one = dh.run_gpu(cuda_lib, model, spec, st, 1, DT)
    ref = dh.run_oracle(model, spec, sub, 1, DT)
    # this state starts with feet pushed well into the ground (stiff contact forces): 5x the free-flight bound
    assert _err(one["q"][pick], ref["q"]) <= 1e-4 and _err(one["pos"][pick], ref["pos"]) <= 1e-4
    assert (np.abs(one["qd"][pick] - ref["qd"]) / np.maximum(1.0, np.abs(ref["qd"]))).max() <= 1e-2
    long = dh.run_gpu(cuda_lib, model, spec, st, 16, DT)
    for k in long:
        assert np.isfinite(long[k]).all(), k
    assert (long["pos"][:, 2] > -0.5).mean() > 0.999  # nobody fell through the ground plane


def test_substeps_in_one_launch_equal_separate_launches(cuda_lib):
    model, spec = dh.specs("Humanoid")
    st = dh.random_state(model, 128, 4)
    a = dh.run_gpu(cuda_lib, model, spec, st, 2, DT, calls=1)
    b = dh.run_gpu(cuda_lib, model, spec, st, 1, DT, calls=2)
    for k in a:
        assert np.array_equal(a[k], b[k]), k


def test_free_flight_momentum_without_gravity(cuda_lib):
    """No gravity, no contact: linear momentum is conserved by the continuous dynamics; the first-order
    integrator drifts by O(dt^2) per step in generalized coordinates.  The kernel must (a) show the SAME
    drift as the float64 oracle (agreement 1e-3 m/s) and (b) stay within the integrator's bound."""
    model, spec = dh.specs("Humanoid")
    model.gravity = (0.0, 0.0, 0.0)
    model.damping[:] = 0
    model.stiffness[:] = 0
    st = dh.random_state(model, 64, 5, height=10.0)
    st["actions"][:] = 0
    com0 = _com_velocity(model, spec, st)
    g = dh.run_gpu(cuda_lib, model, spec, st, 2, DT, calls=16)
    o = dh.run_oracle(model, spec, st, 2, DT, calls=16)
    as_state = lambda r: dict({k: r[k].astype(np.float32) for k in ("pos", "quat", "vel", "q", "qd")}, actions=st["actions"])
    com_g = _com_velocity(model, spec, as_state(g))
    com_o = _com_velocity(model, spec, as_state(o))
    assert np.abs(com_g - com_o).max() < 1e-3
    assert np.abs(com_g - com0).max() < 0.1


def _com_velocity(model, spec, st):
    """Linear momentum / mass via finite differencing the oracle's kinematics is overkill; use the
    oracle integrator with zero gravity for one tiny step and difference the mass-weighted COM."""
    from oracle import dynamics as dyn

    desc = dyn.make_desc(model)
    d = {k: np.ascontiguousarray(st[k].astype(np.float64)) for k in ("pos", "quat", "vel", "q", "qd")}
    return np.stack([_lin_momentum(model, d, i) for i in range(d["pos"].shape[0])]) / model.total_mass()


def _lin_momentum(model, d, i):
    # forward kinematics in numpy (float64): p = sum m_k v_com_k
    from scipy.spatial.transform import Rotation as Rot

    L = model.num_links
    R = [None] * L
    p = [None] * L
    w = [None] * L
    v = [None] * L
    q = d["quat"][i]
    R[0] = Rot.from_quat([q[1], q[2], q[3], q[0]]).as_matrix()
    p[0] = d["pos"][i]
    v[0] = d["vel"][i, :3]
    w[0] = d["vel"][i, 3:]
    mom = np.zeros(3)
    for k in range(L):
        if k > 0:
            pa = model.parent[k]
            jq = model.joint_quat[k]
            Rj = Rot.from_quat([jq[1], jq[2], jq[3], jq[0]]).as_matrix()
            ax = model.joint_axis[k]
            dof = model.dof_index[k]
            R[k] = R[pa] @ Rj @ Rot.from_rotvec(ax * d["q"][i, dof]).as_matrix()
            r = R[pa] @ model.joint_pos[k]
            p[k] = p[pa] + r
            w[k] = w[pa] + R[k] @ ax * d["qd"][i, dof]
            v[k] = v[pa] + np.cross(w[pa], r)
        rc = R[k] @ model.com[k]
        mom += model.mass[k] * (v[k] + np.cross(w[k], rc))
    return mom


def test_non_finite_state_is_parked_for_reset(cuda_lib):
    model, spec = dh.specs("Ant")
    st = dh.random_state(model, 64, 6)
    st["qd"][3, 2] = np.nan
    st["vel"][7, 0] = np.inf
    from omniisaacgymenvs_b200 import _lib

    parked0 = _lib.nonfinite_env_count()
    g = dh.run_gpu(cuda_lib, model, spec, st, 2, DT)
    assert _lib.nonfinite_env_count() == parked0 + 2, "every articulation kernel counts the envs it parks"
    for k in g:
        assert np.isfinite(g[k]).all(), k
    assert g["pos"][3, 2] < spec.termination_height and g["pos"][7, 2] < spec.termination_height
    assert g["pos"][0, 2] > 0.0


@pytest.mark.parametrize("robot", ["Humanoid", "Ant"])
@pytest.mark.parametrize("n", [1, 31, 33, 77])
def test_ragged_env_counts_equal_rows_of_a_full_launch(cuda_lib, robot, n):
    """Env counts that do not fill a warp / a 32-env CTA: the padded lanes of the last group must neither change the
    real rows (bit for bit the rows of a 96-env launch) nor write past the end of the tensors."""
    model, spec = dh.specs(robot)
    st = dh.random_state(model, 96, 17, height=model.spawn_height * 0.8, tilt=0.2)
    full = dh.run_gpu(cuda_lib, model, spec, st, 2, DT)
    part = dh.run_gpu(cuda_lib, model, spec, {k: np.ascontiguousarray(v[:n]) for k, v in st.items()}, 2, DT)
    for k in full:
        assert part[k].shape[0] == n and np.array_equal(part[k], full[k][:n]), k


def test_non_finite_state_in_a_limb_parks_the_env_in_every_kernel(cuda_lib):
    """A NaN that only one limb sees (role-split kernel: only one warp of the CTA) must park the WHOLE env: root below
    the termination height, every joint zeroed, sensors zeroed -- and leave the neighbours alone."""
    model, spec = dh.specs("Humanoid")
    st = dh.random_state(model, 64, 9)
    clean = dh.run_gpu(cuda_lib, model, spec, st, 2, DT)
    for d in (model.num_dof - 1, 0, model.num_dof // 2):  # last / first / middle dof: different limbs and the trunk
        bad = {k: v.copy() for k, v in st.items()}
        bad["qd"][5, d] = np.nan
        g = dh.run_gpu(cuda_lib, model, spec, bad, 2, DT)
        for k in g:
            assert np.isfinite(g[k]).all(), (d, k)
        assert g["pos"][5, 2] < spec.termination_height, d
        assert np.all(g["q"][5] == 0.0) and np.all(g["qd"][5] == 0.0), d
        keep = np.arange(64) != 5
        for k in g:
            assert np.array_equal(g[k][keep], clean[k][keep]), (d, k)


def test_model_create_rejects_malformed_trees(cuda_lib):
    import ctypes as C

    from omniisaacgymenvs_b200 import _lib
    from omniisaacgymenvs_b200.robots.model import fill_desc

    model, _ = dh.specs("Ant")
    desc = _lib.OigeModelDesc()
    fill_desc(model, desc)
    desc.parent[3] = 5  # parent after child
    assert not cuda_lib.oige_model_create(C.byref(desc))
    assert b"malformed" in cuda_lib.oige_last_error()
    fill_desc(model, desc)
    desc.num_links = 40
    assert not cuda_lib.oige_model_create(C.byref(desc))


# ---- prismatic joint + welded base: the Cartpole tree ---------------------------------
class _UnitSpec:
    power_scale = 1.0


def _cartpole_state(n, seed, floating=False):
    rng = np.random.default_rng(seed)
    f32 = lambda a: np.ascontiguousarray(np.asarray(a, np.float64).astype(np.float32))
    pos = np.zeros((n, 3))
    pos[:, :2] = rng.normal(size=(n, 2)) * 3.0
    pos[:, 2] = 50.0 if floating else 2.0
    quat = np.zeros((n, 4))
    quat[:, 0] = 1.0
    vel = np.zeros((n, 6))
    if floating:
        axis = rng.normal(size=(n, 3))
        axis /= np.linalg.norm(axis, axis=1, keepdims=True)
        ang = 0.5 * rng.normal(size=(n, 1))
        quat = np.concatenate([np.cos(ang / 2), np.sin(ang / 2) * axis], axis=1)
        vel = 0.5 * rng.normal(size=(n, 6))
    q = np.stack([rng.uniform(-2.5, 2.5, n), rng.uniform(-3.0, 3.0, n)], axis=1)
    qd = np.stack([rng.normal(size=n), 2.0 * rng.normal(size=n)], axis=1)
    act = np.stack([rng.uniform(-400, 400, n), np.zeros(n)], axis=1)  # efforts: tasks/cartpole.py:104-105
    return {"pos": f32(pos), "quat": f32(quat), "vel": f32(vel), "q": f32(q), "qd": f32(qd), "actions": f32(act)}


@pytest.mark.parametrize("floating", [False, True])
def test_cartpole_single_substep(cuda_lib, floating):
    from omniisaacgymenvs_b200.robots.model import build_model

    model = build_model("Cartpole", fixed_base=not floating)
    st = _cartpole_state(500, 40, floating)
    g = dh.run_gpu(cuda_lib, model, _UnitSpec, st, 1, DT)
    o = dh.run_oracle(model, _UnitSpec, st, 1, DT)
    assert _err(g["q"], o["q"]) < 2e-5
    assert (np.abs(g["qd"] - o["qd"]) / np.maximum(1.0, np.abs(o["qd"]))).max() < 1e-4
    assert _err(g["pos"], o["pos"]) < 2e-5 and _quat_err(g["quat"], o["quat"]) < 2e-5
    assert (np.abs(g["vel"] - o["vel"]) / np.maximum(1.0, np.abs(o["vel"]))).max() < 2e-3
    if not floating:  # the welded rail is untouched, bit for bit
        assert np.array_equal(g["pos"], st["pos"].astype(np.float64)) and np.abs(g["vel"]).max() == 0.0


@pytest.mark.parametrize("floating", [False, True])
def test_cartpole_rollout_120_substeps(cuda_lib, floating):
    from omniisaacgymenvs_b200.robots.model import build_model

    model = build_model("Cartpole", fixed_base=not floating)
    st = _cartpole_state(500, 41, floating)
    st["actions"][:] *= 0.02  # keep the cart inside the rail limits for a second of simulated time
    g = dh.run_gpu(cuda_lib, model, _UnitSpec, st, 2, DT, calls=60)
    o = dh.run_oracle(model, _UnitSpec, st, 2, DT, calls=60)
    assert np.abs(o["q"][:, 1] - st["q"][:, 1]).max() > 1.0, "the poles must actually swing"
    per_env = np.abs(g["q"] - o["q"]).max(axis=1)
    assert np.median(per_env) < 2e-4 and (per_env < 1e-2).mean() >= 0.99  # near-upright poles amplify rounding
    assert _err(g["pos"], o["pos"]) < 5e-3


def test_cartpole_rail_limits_hold_the_cart(cuda_lib):
    """A cart pushed hard against the +-4 m rail limit is stopped by the (implicit) limit spring, as in the oracle."""
    from omniisaacgymenvs_b200.robots.model import build_model

    model = build_model("Cartpole")
    st = _cartpole_state(64, 42)
    st["q"][:, 0] = 3.9
    st["qd"][:, 0] = 3.0
    st["actions"][:, 0] = 50.0
    g = dh.run_gpu(cuda_lib, model, _UnitSpec, st, 2, DT, calls=30)
    o = dh.run_oracle(model, _UnitSpec, st, 2, DT, calls=30)
    assert o["q"][:, 0].max() < 4.3 and g["q"][:, 0].max() < 4.3
    assert (np.abs(g["q"] - o["q"]).max(axis=1) < 2e-2).mean() >= 0.95


def test_model_create_rejects_unknown_joint_type(cuda_lib):
    import ctypes as C

    from omniisaacgymenvs_b200 import _lib
    from omniisaacgymenvs_b200.robots.model import build_model, fill_desc

    desc = _lib.OigeModelDesc()
    fill_desc(build_model("Cartpole"), desc)
    desc.joint_type[1] = 7
    assert not cuda_lib.oige_model_create(C.byref(desc))
    assert b"malformed" in cuda_lib.oige_last_error()


# ---- self-collision (cfg/task/Humanoid.yaml:80 enable_self_collisions: True) ----------------------------------------

def _sc_state(model, n, seed, height=50.0, min_depth=0.01):
    f32 = lambda a: np.ascontiguousarray(a.astype(np.float32))
    s = dh.self_contact_states(model, n, seed, min_depth=min_depth, height=height)
    rng = np.random.default_rng(seed + 1000)
    return {k: f32(v) for k, v in s.items()} | {"actions": f32(rng.uniform(-1, 1, size=s["q"].shape))}


def test_self_collision_single_substep_matches_oracle(cuda_lib, k2_kernel):
    """Every env starts with at least one sphere pair of non-adjacent bodies in contact; the kernel that integrates
    self-collision (K2-F) must agree with the float64 oracle, and differ from the run without it."""
    from omniisaacgymenvs_b200.robots.model import build_model

    model = build_model("Humanoid")
    spec = dh.lo.SPECS["Humanoid"]
    st = _sc_state(model, 256, 3)
    if k2_kernel in ("warp", "thread"):
        with pytest.raises(RuntimeError, match="self-collision"):
            dh.run_gpu(cuda_lib, model, spec, st, 1, DT)
        return
    g = dh.run_gpu(cuda_lib, model, spec, st, 1, DT)
    o = dh.run_oracle(model, spec, st, 1, DT)
    off = dh.run_gpu(cuda_lib, build_model("Humanoid", enable_self_collisions=False), spec, st, 1, DT)
    assert np.abs(g["qd"] - off["qd"]).max(axis=1).min() > 1e-2, "the pairs must actually push in every env"
    assert _err(g["q"], o["q"]) < 2e-5 and _err(g["pos"], o["pos"]) < 2e-5 and _quat_err(g["quat"], o["quat"]) < 2e-5
    assert (np.abs(g["qd"] - o["qd"]) / np.maximum(1.0, np.abs(o["qd"]))).max() < 2e-3
    assert (np.abs(g["vel"] - o["vel"]) / np.maximum(1.0, np.abs(o["vel"]))).max() < 2e-3


def test_self_collision_rollout_separates_limbs_like_the_oracle(cuda_lib, k2_kernel):
    """32 substeps from interpenetrating limbs (free flight, no actuation): the limbs are pushed apart, the joint
    trajectories follow the oracle, and the root's momentum change stays at rounding level (internal forces)."""
    if k2_kernel in ("warp", "thread"):
        pytest.skip("only K2-F integrates self-collision")
    from omniisaacgymenvs_b200.robots.model import build_model

    model = build_model("Humanoid", gravity=(0.0, 0.0, 0.0))
    spec = dh.lo.SPECS["Humanoid"]
    st = _sc_state(model, 128, 5, min_depth=0.02)
    st["actions"][:] = 0.0
    st["qd"][:] = 0.0
    d0 = np.array([dh.self_contact_depths(model, st, e).max() for e in range(128)])
    g = dh.run_gpu(cuda_lib, model, spec, st, 32, DT)
    o = dh.run_oracle(model, spec, st, 32, DT)
    d1 = np.array([dh.self_contact_depths(model, g, e).max() for e in range(128)])
    assert np.isfinite(g["q"]).all() and np.median(d1) < 0.5 * np.median(d0), (np.median(d0), np.median(d1))
    per_env = np.abs(g["q"] - o["q"]).max(axis=1)
    assert np.median(per_env) < 2e-4 and (per_env < 2e-2).mean() >= 0.97, (np.median(per_env), per_env.max())


def test_self_collision_in_the_full_step_keeps_limbs_apart(cuda_lib, k2_kernel):
    """Random-policy episodes through the public step: with self-collision the limbs' capsules overlap far less
    (penalty spring: a few centimetres under full motor torque) than without it, where limbs pass through the body."""
    if k2_kernel != "role":
        pytest.skip("whole-step check, once")
    import torch

    from omniisaacgymenvs_b200.utils.task_util import make_env
    from omniisaacgymenvs_b200.robots.model import build_model

    model = build_model("Humanoid")
    overlap = {}
    for on in (True, False):
        env, task = make_env("Humanoid", num_envs=256, overrides={} if on else {"task.sim.Humanoid.enable_self_collisions": False})
        env.reset()
        g = torch.Generator(device="cuda").manual_seed(0)
        total = 0.0
        for t in range(60):
            env.step(torch.rand(256, 21, device="cuda", generator=g) * 2 - 1)
            if t % 6 == 5:
                v = task.get_robot()
                pos, quat = v.get_world_poses(clone=True)
                s = {"pos": pos.cpu().numpy().astype(np.float64), "quat": quat.cpu().numpy().astype(np.float64),
                     "q": v.get_joint_positions(clone=True).cpu().numpy().astype(np.float64)}
                total += sum(np.clip(dh.self_contact_depths(model, s, e), 0.0, None).sum() for e in range(0, 256, 4))
        overlap[on] = total
    print("summed capsule overlap (m): with self-collision", overlap[True], "without", overlap[False])
    assert overlap[True] < 0.5 * overlap[False], overlap
