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
