"""CPU: sanity of the float64 dynamics oracle itself (energy, momentum, rest), so that a GPU/oracle
agreement means something."""
import numpy as np
import pytest

from omniisaacgymenvs_b200.robots.model import build_model
from oracle import dynamics as dyn


def _free(model):
    model.damping[:] = 0
    model.stiffness[:] = 0
    model.limit_lower[:] = -1e9
    model.limit_upper[:] = 1e9
    model.angular_damping = 0.0          # conservative system: no body damping, no velocity clamps, no limb contacts
    model.max_angular_velocity = 0.0
    model.enable_self_collisions = False
    return model


@pytest.mark.parametrize("robot", ["Humanoid", "Ant"])
def test_energy_drift_shrinks_with_dt(robot):
    m = _free(build_model(robot))
    desc = dyn.make_desc(m)
    rng = np.random.default_rng(0)
    D = m.num_dof
    base = dict(pos=np.array([[0, 0, 50.0]]), quat=np.array([[1.0, 0, 0, 0]]), vel=rng.normal(size=(1, 6)),
                q=rng.uniform(-0.3, 0.3, (1, D)), qd=rng.normal(size=(1, D)))
    e0 = dyn.energy(desc, *[base[k] for k in ("pos", "quat", "vel", "q", "qd")])
    drift = []
    for dt in (0.0083, 0.00083):
        s = {k: v.copy() for k, v in base.items()}
        assert dyn.step(desc, s["pos"], s["quat"], s["vel"], s["q"], s["qd"], np.zeros((1, D)), None,
                        int(round(0.5 / dt)), dt) == 0
        drift.append(abs(dyn.energy(desc, *[s[k] for k in ("pos", "quat", "vel", "q", "qd")]) - e0) / abs(e0))
    assert drift[0] < 2e-3 and drift[1] < drift[0] / 5  # first-order integrator


def test_free_fall_matches_closed_form():
    m = _free(build_model("Ant"))
    desc = dyn.make_desc(m)
    D = m.num_dof
    s = dict(pos=np.array([[0, 0, 100.0]]), quat=np.array([[1.0, 0, 0, 0]]), vel=np.zeros((1, 6)),
             q=np.zeros((1, D)), qd=np.zeros((1, D)))
    n, dt = 100, 0.0083
    g = float(np.float32(9.81))  # the model struct stores gravity in fp32
    dyn.step(desc, s["pos"], s["quat"], s["vel"], s["q"], s["qd"], np.zeros((1, D)), None, n, dt)
    assert abs(s["vel"][0, 2] - (-g * n * dt)) < 1e-9
    # semi-implicit Euler: z = z0 - g dt^2 n(n+1)/2
    assert abs(s["pos"][0, 2] - (100.0 - g * dt * dt * n * (n + 1) / 2)) < 1e-9
    assert np.abs(s["q"]).max() < 1e-9


def test_ant_comes_to_rest_on_the_ground():
    m = build_model("Ant")
    desc = dyn.make_desc(m)
    D = m.num_dof
    lim = m.dof_limits()
    s = dict(pos=np.array([[0, 0, m.spawn_height]]), quat=np.array([[1.0, 0, 0, 0]]), vel=np.zeros((1, 6)),
             q=np.clip(np.zeros((1, D)), lim[:, 0], lim[:, 1]).astype(np.float64), qd=np.zeros((1, D)))
    sens = np.zeros((1, m.num_sensors, 6))
    assert dyn.step(desc, s["pos"], s["quat"], s["vel"], s["q"], s["qd"], np.zeros((1, D)), sens, 240, 0.0083) == 0
    assert np.abs(s["vel"]).max() < 1e-2 and np.abs(s["qd"]).max() < 1e-2
    assert 0.2 < s["pos"][0, 2] < m.spawn_height
    assert sens[0, :, 2].min() >= 0.0  # the ground only pushes


def test_openmp_threads_do_not_change_results():
    m = build_model("Humanoid")
    desc = dyn.make_desc(m)
    rng = np.random.default_rng(1)
    n, D = 64, m.num_dof
    mk = lambda: dict(pos=np.tile([[0, 0, 1.3]], (n, 1)).astype(np.float64), quat=np.tile([[1.0, 0, 0, 0]], (n, 1)),
                      vel=np.zeros((n, 6)), q=np.zeros((n, D)), qd=np.zeros((n, D)))
    eff = rng.uniform(-20, 20, (n, D))
    a, b = mk(), mk()
    dyn.step(desc, a["pos"], a["quat"], a["vel"], a["q"], a["qd"], eff, None, 8, 0.0083, num_threads=1)
    dyn.step(desc, b["pos"], b["quat"], b["vel"], b["q"], b["qd"], eff, None, 8, 0.0083, num_threads=4)
    for k in a:
        assert np.array_equal(a[k], b[k])


# ---- prismatic joints and a welded base (Cartpole) ---------------------------------
def _cartpole_accel(m, x, th, xd, thd, force):
    """Textbook cart-pole equations (Lagrange), cart along +y, pole hinge about +x, theta = 0 upright."""
    M, mp = m.mass[1], m.mass[2]
    l = m.com[2, 2]
    I = m.inertia[2, 0]  # about the hinge axis (x), at the centre of mass
    g = float(np.float32(9.81))
    # (M+m) xdd - m l cos(th) thdd + m l sin(th) thd^2 = F ;  (I + m l^2) thdd - m l cos(th) xdd - m g l sin(th) = 0
    A = np.array([[M + mp, -mp * l * np.cos(th)], [-mp * l * np.cos(th), I + mp * l * l]])
    b = np.array([force - mp * l * np.sin(th) * thd * thd, mp * g * l * np.sin(th)])
    return np.linalg.solve(A, b)


def test_cartpole_accelerations_match_the_textbook_equations():
    m = build_model("Cartpole", angular_damping=0.0)
    assert m.fixed_base and list(m.joint_type) == [0, 1, 0] and m.dof_names == ["cartJoint", "poleJoint"]
    desc = dyn.make_desc(m)
    rng = np.random.default_rng(3)
    dt = 1e-3
    for _ in range(20):
        x, th = rng.uniform(-2, 2), rng.uniform(-3, 3)
        xd, thd = rng.normal(), 2 * rng.normal()
        f = rng.uniform(-400, 400)
        s = dict(pos=np.array([[1.0, -2.0, 2.0]]), quat=np.array([[1.0, 0, 0, 0]]), vel=np.zeros((1, 6)),
                 q=np.array([[x, th]]), qd=np.array([[xd, thd]]))
        assert dyn.step(desc, s["pos"], s["quat"], s["vel"], s["q"], s["qd"], np.array([[f, 0.0]]), None, 1, dt) == 0
        acc = (s["qd"][0] - np.array([xd, thd])) / dt
        np.testing.assert_allclose(acc, _cartpole_accel(m, x, th, xd, thd, f), rtol=1e-9, atol=1e-9)
        # the welded rail did not move
        assert np.array_equal(s["pos"], [[1.0, -2.0, 2.0]]) and np.array_equal(s["vel"], np.zeros((1, 6)))


def test_cartpole_energy_is_conserved_without_actuation():
    m = build_model("Cartpole", angular_damping=0.0)
    desc = dyn.make_desc(m)
    base = dict(pos=np.array([[0, 0, 2.0]]), quat=np.array([[1.0, 0, 0, 0]]), vel=np.zeros((1, 6)),
                q=np.array([[0.3, 0.4]]), qd=np.array([[-0.5, 1.0]]))
    e0 = dyn.energy(desc, *[base[k] for k in ("pos", "quat", "vel", "q", "qd")])
    drift = []
    for dt in (1e-3, 1e-4):
        s = {k: v.copy() for k, v in base.items()}
        assert dyn.step(desc, s["pos"], s["quat"], s["vel"], s["q"], s["qd"], np.zeros((1, 2)), None,
                        int(round(1.0 / dt)), dt) == 0
        drift.append(abs(dyn.energy(desc, *[s[k] for k in ("pos", "quat", "vel", "q", "qd")]) - e0) / abs(e0))
    assert drift[0] < 5e-3 and drift[1] < drift[0] / 5


def test_prismatic_joint_on_a_floating_base_conserves_momentum():
    """A tumbling, free-floating cart-pole (no gravity): linear momentum is exactly conserved by the
    integrator, and the energy drift shrinks with dt -- exercises the Coriolis terms of the slider."""
    m = build_model("Cartpole", fixed_base=False, gravity=(0.0, 0.0, 0.0), angular_damping=0.0)
    desc = dyn.make_desc(m)
    rng = np.random.default_rng(5)
    base = dict(pos=np.array([[0, 0, 50.0]]), quat=np.array([[1.0, 0, 0, 0]]), vel=rng.normal(size=(1, 6)),
                q=np.array([[0.5, 0.7]]), qd=np.array([[1.5, -2.0]]))

    def momentum(s):
        # finite-difference the centre of mass: total mass x com velocity, from two energy-free evaluations
        from dyn_helpers import link_coms
        return link_coms(m, s)

    e0 = dyn.energy(desc, *[base[k] for k in ("pos", "quat", "vel", "q", "qd")])
    drift = []
    for dt in (1e-3, 1e-4):
        s = {k: v.copy() for k, v in base.items()}
        c0 = momentum(s)
        n = int(round(0.5 / dt))
        assert dyn.step(desc, s["pos"], s["quat"], s["vel"], s["q"], s["qd"], np.zeros((1, 2)), None, 1, dt) == 0
        c1 = momentum(s)
        assert dyn.step(desc, s["pos"], s["quat"], s["vel"], s["q"], s["qd"], np.zeros((1, 2)), None, n, dt) == 0
        c2 = momentum(s)
        assert dyn.step(desc, s["pos"], s["quat"], s["vel"], s["q"], s["qd"], np.zeros((1, 2)), None, 1, dt) == 0
        c3 = momentum(s)
        v_start, v_end = (c1 - c0) / dt, (c3 - c2) / dt
        np.testing.assert_allclose(v_end, v_start, atol=5e-3 * max(1.0, np.abs(v_start).max()) * (dt / 1e-3))
        drift.append(abs(dyn.energy(desc, *[s[k] for k in ("pos", "quat", "vel", "q", "qd")]) - e0) / abs(e0))
    assert drift[1] < drift[0] / 5


def _self_contact_states(m, n, seed, min_depth=0.01):
    import dyn_helpers as dh

    return dh.self_contact_states(m, n, seed, min_depth)


def test_self_collision_pairs_follow_the_body_tree():
    """PhysX semantics: every pair of bodies with shapes except parent / child (cfg/task/Humanoid.yaml:80)."""
    m = build_model("Humanoid")
    assert m.enable_self_collisions and len(m.collision_pairs) == 13 * 12 // 2 - 12
    names = {(m.link_names[a], m.link_names[b]) for a, b in m.collision_pairs}
    assert ("right_elbow", "left_elbow") in names and ("torso", "right_knee") in names
    assert ("torso", "abdomen_y") not in names and ("right_hip_y", "right_knee") not in names  # jointed bodies
    ant = build_model("Ant")
    assert not ant.enable_self_collisions and dyn.make_desc(ant).num_collision_pairs == 0  # Ant.yaml:82
    assert dyn.make_desc(m).num_collision_pairs == 66
    assert dyn.make_desc(build_model("Humanoid", enable_self_collisions=False)).num_collision_pairs == 0


def test_self_collision_is_an_internal_force():
    """Limbs in contact are pushed apart (the joint velocities change at O(dt)) while the centre of mass moves as it
    does without self-collision (difference O(dt^3)): the pair forces are equal and opposite (no gravity, free flight)."""
    import dyn_helpers as dh

    kw = dict(gravity=(0.0, 0.0, 0.0), angular_damping=0.0)
    on, off = build_model("Humanoid", **kw), build_model("Humanoid", enable_self_collisions=False, **kw)
    base = _self_contact_states(on, 8, 1)
    D = on.num_dof
    res = {}
    for name, m in (("on", on), ("off", off)):
        s = {k: v.copy() for k, v in base.items()}
        assert dyn.step(dyn.make_desc(m), s["pos"], s["quat"], s["vel"], s["q"], s["qd"], np.zeros((8, D)), None, 1, 1e-3) == 0
        res[name] = s
    assert np.abs(res["on"]["qd"] - res["off"]["qd"]).max(axis=1).min() > 1e-2, "every env had a pair in contact"
    for e in range(8):
        com = [dh.link_coms(on, {k: v[e:e + 1] for k, v in res[n_].items()}) for n_ in ("on", "off")]
        np.testing.assert_allclose(com[0], com[1], atol=1e-7)


def test_self_collision_separates_interpenetrating_limbs():
    """Held only by the joint springs, limbs that start in contact end up (nearly) out of contact."""
    import dyn_helpers as dh

    m = build_model("Humanoid", gravity=(0.0, 0.0, 0.0))
    s = _self_contact_states(m, 6, 2, min_depth=0.02)
    s["qd"][:] = 0.0
    d0 = np.array([dh.self_contact_depths(m, s, e).max() for e in range(6)])
    assert dyn.step(dyn.make_desc(m), s["pos"], s["quat"], s["vel"], s["q"], s["qd"], np.zeros((6, m.num_dof)), None, 240, 0.0083) == 0
    d1 = np.array([dh.self_contact_depths(m, s, e).max() for e in range(6)])
    assert np.isfinite(s["q"]).all() and (d1 < 0.5 * d0).all(), (d0, d1)
