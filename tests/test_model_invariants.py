"""Model-level checks of the articulation step that do NOT go through the float64 oracle (ADVICE
item 1): weight on the foot sensors of a resting Ant, vertical impulse / momentum balance of a Humanoid on its feet,
friction cone of the sensed foot forces, total mass / centre of mass against values recomputed by hand from the MJCF
geoms, and the random-policy episode-length band.  The reference pins none of this (PhysX is closed); these are the
invariants any rigid-body model of the robots must satisfy."""
import math

import numpy as np
import pytest
import torch

from omniisaacgymenvs_b200.robots import mjcf_tables
from omniisaacgymenvs_b200.robots.model import build_model
from oracle import locomotion_oracle as lo
from dyn_helpers import link_coms, link_frames, run_gpu

G = 9.81
DT = 1.0 / 120.0


# ---------------------------------------------------------------------------------------------------------------
# CPU: masses and centre of mass recomputed by hand from the geoms (capsule = cylinder + two half spheres)
# ---------------------------------------------------------------------------------------------------------------
def _geom_mass_com(g, density):
    if g[0] == "sphere":
        return density * 4.0 / 3.0 * math.pi * g[2] ** 3, np.array(g[1], float)
    a, b, r = np.array(g[1], float), np.array(g[2], float), g[3]
    length = float(np.linalg.norm(b - a))
    return density * (math.pi * r * r * length + 4.0 / 3.0 * math.pi * r ** 3), 0.5 * (a + b)


@pytest.mark.parametrize("robot,table", [("Humanoid", mjcf_tables.HUMANOID), ("Ant", mjcf_tables.ANT)])
def test_mass_and_com_equal_hand_computed_mjcf_values(robot, table):
    m = build_model(robot)
    bodies = table["bodies"]()
    # world position of every body at q = 0 with the root at the origin: body quats are identity up to the 0.002 rad
    # tilt of the waist (moves a point 0.4 m away by < 1 mm), so positions add
    pos, total, moment = {}, 0.0, np.zeros(3)
    for b in bodies:
        pos[b["name"]] = np.array(b["pos"], float) + (pos[b["parent"]] if b["parent"] else 0.0)
        for g in b["geoms"]:
            if g[0] not in ("capsule", "sphere"):
                continue
            mg, cg = _geom_mass_com(g, table["density"])
            total += mg
            moment += mg * (pos[b["name"]] + cg)
    assert abs(m.total_mass() - total) < 1e-3 * total
    s = dict(pos=np.zeros((1, 3)), quat=np.array([[1.0, 0, 0, 0]]), q=np.zeros((1, m.num_dof)))
    com = link_coms(m, s)
    assert np.abs(com - moment / total).max() < 2e-3
    # the figures themselves, so that a change of the tables shows up here: nv_humanoid.xml at density 1000 and
    # nv_ant.xml at density 5
    if robot == "Humanoid":
        assert 40.0 < total < 41.7 and -0.46 < com[2] < -0.40  # 40.8 kg; COM 0.43 m below the torso origin (0.91 m at spawn)
    else:
        assert 0.5 < total < 1.2 and abs(com[2]) < 0.05


# ---------------------------------------------------------------------------------------------------------------
# GPU
# ---------------------------------------------------------------------------------------------------------------
def _world_sensor_forces(m, out, e):
    """Ground force on each sensor link of env e in world axes (the kernel reports it in the link frame)."""
    R, _ = link_frames(m, out, e)
    return np.array([R[int(l)] @ out["sensors"][e, k, :3] for k, l in enumerate(m.sensor_link)])  # [force, torque]


@pytest.mark.gpu
def test_resting_ant_foot_sensors_carry_its_weight(cuda_lib):
    m, spec = build_model("Ant"), lo.SPECS["Ant"]
    n = 8
    st = {"pos": np.tile([[0.0, 0.0, 0.55]], (n, 1)), "quat": np.tile([[1.0, 0, 0, 0]], (n, 1)), "vel": np.zeros((n, 6)),
          "q": np.zeros((n, m.num_dof)), "qd": np.zeros((n, m.num_dof)), "actions": np.zeros((n, m.num_dof))}
    lim = m.dof_limits()
    st["q"] = np.clip(st["q"], lim[:, 0], lim[:, 1])
    st = {k: np.ascontiguousarray(v.astype(np.float32)) for k, v in st.items()}
    out = run_gpu(cuda_lib, m, spec, st, 2, DT, calls=360)  # 6 s: settled
    assert np.abs(out["vel"]).max() < 2e-2 and np.abs(out["qd"]).max() < 5e-2, "the ant must be at rest"
    for e in range(n):
        f = _world_sensor_forces(m, out, e)
        assert abs(f[:, 2].sum() - m.total_mass() * G) < 0.01 * m.total_mass() * G
        assert (f[:, 2] > 0).all()
        assert (np.linalg.norm(f[:, :2], axis=1) <= m.friction_mu * f[:, 2] * 1.001 + 1e-4).all()  # friction cone


@pytest.mark.gpu
def test_humanoid_on_its_feet_vertical_impulse_equals_momentum_change(cuda_lib):
    """sum_t (F_z(t) - m g) dt = m (v_com,z(end) - v_com,z(start)) while only the feet touch the ground; the COM velocity
    of the semi-implicit integrator is the backward difference of the COM position."""
    m, spec = build_model("Humanoid"), lo.SPECS["Humanoid"]
    n, steps = 4, 24
    rng = np.random.default_rng(3)
    st = {"pos": np.tile([[0.0, 0.0, 1.34]], (n, 1)), "quat": np.tile([[1.0, 0, 0, 0]], (n, 1)), "vel": np.zeros((n, 6)),
          "q": 0.02 * rng.normal(size=(n, m.num_dof)), "qd": np.zeros((n, m.num_dof)),
          "actions": 0.1 * rng.uniform(-1, 1, size=(n, m.num_dof))}
    st = {k: np.ascontiguousarray(v.astype(np.float32)) for k, v in st.items()}
    mass = m.total_mass()
    coms = [[link_coms(m, {k: v[e:e + 1].astype(np.float64) for k, v in st.items()}) for e in range(n)]]
    impulse = np.zeros(n)
    first_impulse = np.zeros(n)
    touched = np.zeros(n, bool)
    cur = st
    for t in range(steps):
        out = run_gpu(cuda_lib, m, spec, cur, 1, DT)
        coms.append([link_coms(m, {k: out[k][e:e + 1] for k in ("pos", "quat", "q")}) for e in range(n)])
        fz = np.array([_world_sensor_forces(m, out, e)[:, 2].sum() for e in range(n)])
        touched |= fz > 1.0
        if t == 0:
            first_impulse = (fz - mass * G) * DT
        else:
            impulse += (fz - mass * G) * DT
        cur = {k: np.ascontiguousarray(out[k].astype(np.float32)) for k in ("pos", "quat", "vel", "q", "qd")}
        cur["actions"] = st["actions"]
    assert touched.all(), "the feet must reach the ground in this window"
    coms = np.array(coms)  # [steps + 1, n, 3]
    v_start = (coms[1, :, 2] - coms[0, :, 2]) / DT  # COM velocity after substep 0
    v_end = (coms[-1, :, 2] - coms[-2, :, 2]) / DT
    dp = mass * (v_end - v_start)
    scale = mass * G * DT * (steps - 1)
    assert np.abs(impulse - dp).max() < 0.02 * scale, (impulse, dp, first_impulse)


@pytest.mark.gpu
def test_random_policy_episode_length_bands(cuda_lib):
    """A uniform random policy drops the Humanoid within a few dozen control steps (the published Isaac Gym / OIGE
    random-policy episodes are 20-40 steps long); the Ant survives several times longer."""
    from omniisaacgymenvs_b200.utils.task_util import make_env

    mean_len, rate = {}, {}
    for robot, n_act in (("Humanoid", 21), ("Ant", 8)):
        env, task = make_env(robot, num_envs=2048, seed=1)
        env.reset()
        task.read_episode_stats()
        for _ in range(400):
            env.step(torch.rand(2048, n_act, device=task.device) * 2 - 1)
        st = task.read_episode_stats().cpu().numpy()
        rate[robot] = st[1] / st[3]  # episodes finished per env-step
        mean_len[robot] = st[2] / max(st[1], 1.0)
    assert rate["Humanoid"] > 1.0 / 60.0 and 12.0 < mean_len["Humanoid"] < 60.0, (mean_len, rate)
    assert rate["Ant"] < rate["Humanoid"] / 1.5, (mean_len, rate)  # the four-legged Ant rarely tips below 0.31 m


@pytest.mark.gpu
def test_ppo_return_rises_on_the_ant(cuda_lib):
    """Behaviour-level plausibility: a minimal PPO through VecEnvRLGames.step learns to move the Ant --
    the mean reward per step rises from about zero to above 1 within 60 iterations (1 M env-steps)."""
    import os
    import sys

    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import ppo_smoke

    curve = ppo_smoke.run("Ant", iterations=60, num_envs=1024, verbose=False)
    first = sum(c[0] for c in curve[:5]) / 5
    last = sum(c[0] for c in curve[-10:]) / 10
    assert last > 1.0 and last > first + 0.5, (first, last)
