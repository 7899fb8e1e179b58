"""CPU: config resolver and the robot model tables against everything the reference pins."""
import math
import os

import numpy as np
import pytest

from omniisaacgymenvs_b200.articulations import grid_env_origins
from omniisaacgymenvs_b200.robots.model import build_model
from omniisaacgymenvs_b200.utils.config import SimConfig, load_config
from oracle import locomotion_oracle as lo

REF_CFG = "/root/reference/omniisaacgymenvs/cfg/task"


def test_config_defaults_and_overrides():
    c = load_config("Humanoid")
    assert c["task_name"] == "Humanoid" and c["task"]["env"]["numEnvs"] == 4096
    assert c["task"]["sim"]["use_gpu_pipeline"] is True and c["task"]["sim"]["dt"] == 0.0083
    c = load_config("Ant", {"num_envs": 128, "task.env.episodeLength": 50})
    assert c["task"]["env"]["numEnvs"] == 128 and c["task"]["env"]["episodeLength"] == 50
    sc = SimConfig(c)
    assert sc.config["sim_device"] == "cuda:0" and sc.task_config["name"] == "Ant"


@pytest.mark.parametrize("robot", ["Humanoid", "Ant"])
def test_task_constants_equal_oracle_spec(robot):
    env = load_config(robot)["task"]["env"]
    spec = lo.SPECS[robot]
    pairs = [("dofVelocityScale", "dof_vel_scale"), ("angularVelocityScale", "angular_velocity_scale"),
             ("contactForceScale", "contact_force_scale"), ("powerScale", "power_scale"),
             ("headingWeight", "heading_weight"), ("upWeight", "up_weight"), ("actionsCost", "actions_cost_scale"),
             ("energyCost", "energy_cost_scale"), ("jointsAtLimitCost", "joints_at_limit_cost_scale"),
             ("deathCost", "death_cost"), ("terminationHeight", "termination_height"),
             ("alive_reward_scale", "alive_reward_scale"), ("episodeLength", "max_episode_length")]
    for a, b in pairs:
        assert env[a] == getattr(spec, b), a
    assert env["controlFrequencyInv"] == 2 and env["clipActions"] == 1.0 and env["envSpacing"] == 5


@pytest.mark.skipif(not os.path.isdir(REF_CFG), reason="reference tree only exists in the build container")
@pytest.mark.parametrize("robot", ["Humanoid", "Ant"])
def test_our_yaml_env_block_equals_reference_yaml(robot):
    ours = load_config(robot)["task"]
    ref = load_config(robot, task_yaml=os.path.join(REF_CFG, f"{robot}.yaml"))["task"]
    assert ours["env"] == ref["env"]
    for k in ("dt", "gravity", "add_ground_plane", "default_physics_material", "use_gpu_pipeline"):
        assert ours["sim"][k] == ref["sim"][k], k


def test_humanoid_model_matches_what_the_reference_pins():
    m = build_model("Humanoid")
    assert m.num_dof == 21 and m.num_links == 22 and m.num_sensors == 2
    assert 12 + 3 * m.num_dof + 6 * m.num_sensors == 87
    assert list(m.joint_gears) == lo.HUMANOID_GEARS  # tasks/humanoid.py:82-107, in BFS order
    bodies = ["lower_waist"] * 2 + ["right_upper_arm"] * 2 + ["left_upper_arm"] * 2 + ["pelvis", "right_lower_arm",
              "left_lower_arm"] + ["right_thigh"] * 3 + ["left_thigh"] * 3 + ["right_shin", "left_shin"] + \
             ["right_foot"] * 2 + ["left_foot"] * 2
    assert m.dof_body_names == bodies  # the comments of tasks/humanoid.py:84-104
    assert m.spawn_height == 1.34
    assert 35.0 < m.total_mass() < 45.0
    assert sorted(m.dof_index[1:].tolist()) == list(range(21))
    assert all(m.parent[i] < i for i in range(1, m.num_links))


def test_ant_model_matches_what_the_reference_pins():
    m = build_model("Ant")
    assert m.num_dof == 8 and m.num_links == 9 and m.num_sensors == 4
    assert 12 + 3 * m.num_dof + 6 * m.num_sensors == 60
    assert list(m.joint_gears) == [15.0] * 8 and m.spawn_height == 0.5
    lim = np.degrees(m.dof_limits())
    assert np.allclose(lim[:4], [[-40, 40]] * 4, atol=1e-4)


@pytest.mark.parametrize("robot", ["Humanoid", "Ant"])
def test_inertias_are_physical(robot):
    m = build_model(robot)
    for i in range(m.num_links):
        if m.mass[i] == 0:
            assert np.all(m.inertia[i] == 0)
            continue
        xx, yy, zz, xy, xz, yz = m.inertia[i]
        I = np.array([[xx, xy, xz], [xy, yy, yz], [xz, yz, zz]])
        ev = np.linalg.eigvalsh(I)
        assert ev.min() > 0 and ev[0] + ev[1] >= ev[2] * (1 - 1e-9)  # triangle inequality
    # every chain ends in a link with mass (ABA needs a positive joint-space inertia)
    has_child = set(m.parent[1:].tolist())
    for i in range(1, m.num_links):
        if i not in has_child:
            assert m.mass[i] > 0


def test_capsule_inertia_limits():
    from omniisaacgymenvs_b200.robots.model import _geom_inertial

    m, c, I = _geom_inertial(("sphere", (0, 0, 0), 0.1), 1000.0)
    assert abs(m - 1000 * 4 / 3 * math.pi * 1e-3) < 1e-9 and abs(I[0, 0] - 0.4 * m * 0.01) < 1e-12
    # zero-length capsule == sphere
    m2, _, I2 = _geom_inertial(("capsule", (0, 0, 0), (0, 0, 1e-12), 0.1), 1000.0)
    assert abs(m2 - m) < 1e-6 and abs(I2[0, 0] - I[0, 0]) < 1e-7
    # long thin capsule -> rod m L^2 / 12
    m3, _, I3 = _geom_inertial(("capsule", (0, 0, -5), (0, 0, 5), 0.001), 1000.0)
    assert abs(I3[0, 0] / (m3 * 100 / 12) - 1) < 1e-3


def test_grid_origins_layout_and_sharding():
    g = grid_env_origins(4096, 5.0)
    assert g.shape == (4096, 3) and np.all(g[:, 2] == 0)
    assert abs(g[:, 0].mean()) < 1e-3 and abs(g[:, 1].mean()) < 1e-3
    d = np.linalg.norm(g[1] - g[0])
    assert abs(d - 5.0) < 1e-6
    assert len({(round(x, 3), round(y, 3)) for x, y, _ in g}) == 4096
    part = grid_env_origins(4096, 5.0, offset=1024, count=512)
    assert np.array_equal(part, g[1024:1536])
