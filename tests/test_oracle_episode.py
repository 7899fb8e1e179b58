"""CPU: the oracle restatement of the HOST-SIDE rows (reset_idx, efforts, post-physics ordering) against
episodes recorded from the VERBATIM reference task classes (oracle/ref_task_driver.py -> tests/golden/
episode_*.npz: post_reset, then 5 x [pre_physics_step, injected state, progress += 1, get_observations,
calculate_metrics, is_done] through tasks/shared/locomotion.py:80-183, humanoid.py:81-127, ant.py:79-95)."""
import os

import numpy as np
import pytest
import torch

import helpers
from oracle import locomotion_oracle as lo
from oracle import ref_loader


def _load(robot):
    z = np.load(os.path.join(helpers.GOLDEN_DIR, f"episode_{robot.lower()}.npz"))
    return {k: torch.from_numpy(z[k]) for k in z.files}


@pytest.mark.parametrize("robot", ["Humanoid", "Ant"])
def test_oracle_reproduces_reference_episode(robot):
    torch.set_num_threads(1)
    ep, spec = _load(robot), lo.SPECS[robot]
    n, steps = ep["actions"].shape[1], ep["actions"].shape[0]
    # post_reset constants (locomotion.py:147-171; humanoid.py:82-112; ant.py:80-84)
    assert torch.equal(ep["joint_gears"], torch.tensor(spec.joint_gears))
    assert torch.equal(ep["motor_effort_ratio"], spec.motor_effort_ratio())
    assert torch.equal(ep["initial_root_pos"], ep["spawn_pos"])
    one = lambda v: torch.tensor(v, dtype=torch.float32).repeat(n, 1)
    s = {"pos": ep["spawn_pos"].clone(), "quat": ep["initial_root_rot"].clone(), "vel": torch.zeros(n, 6),
         "dof_pos": torch.zeros(n, spec.num_dof), "dof_vel": torch.zeros(n, spec.num_dof),
         "potentials": torch.full((n,), -1000.0 / spec.dt), "prev_potentials": torch.full((n,), -1000.0 / spec.dt),
         "reset_buf": torch.ones(n, dtype=torch.int64), "progress_buf": torch.zeros(n, dtype=torch.int64),
         "initial_root_pos": ep["initial_root_pos"], "initial_root_rot": ep["initial_root_rot"],
         "initial_dof_pos": ep["initial_dof_pos"], "targets": one([1000.0, 0.0, 0.0]),
         "inv_start_rot": one([1.0, 0.0, 0.0, 0.0]), "basis_vec0": one([1.0, 0.0, 0.0]), "basis_vec1": one([0.0, 0.0, 1.0]),
         "dof_lo": ep["dof_limits_lower"], "dof_hi": ep["dof_limits_upper"]}
    lo.reset_idx(spec, s, torch.arange(n), ep["post_reset_rand_pos"], ep["post_reset_rand_vel"])
    assert torch.equal(s["dof_pos"], ep["post_reset_dof_pos"]) and torch.equal(s["dof_vel"], ep["post_reset_dof_vel"])
    assert torch.equal(s["potentials"], ep["post_reset_potentials"])
    for t in range(steps):
        ids = ep["reset_ids"][t].nonzero(as_tuple=False).squeeze(-1)
        assert torch.equal(ids, s["reset_buf"].nonzero(as_tuple=False).squeeze(-1)), "reset sets must agree"
        if len(ids):
            lo.reset_idx(spec, s, ids, ep["rand_pos"][t][ids], ep["rand_vel"][t][ids])
        for k, g in (("dof_pos", "after_reset_dof_pos"), ("dof_vel", "after_reset_dof_vel"), ("pos", "after_reset_pos"),
                     ("quat", "after_reset_quat"), ("vel", "after_reset_vel"), ("potentials", "after_reset_potentials")):
            assert torch.equal(s[k], ep[g][t]), f"step {t}: {k} after reset"
        assert torch.equal(lo.joint_efforts(spec, ep["actions"][t]), ep["efforts"][t])
        s.update(pos=ep["inj_pos"][t].clone(), quat=ep["inj_quat"][t].clone(), vel=ep["inj_vel"][t].clone(),
                 dof_pos=ep["inj_dof_pos"][t].clone(), dof_vel=ep["inj_dof_vel"][t].clone(),
                 sensors=ep["inj_sensors"][t].clone(), actions=ep["actions"][t].clone())
        if t == int(ep["progress_override_step"]):
            s["progress_buf"][: n // 4] = spec.max_episode_length - 2
        out = lo.post_physics_step(spec, s)
        for k in ("reset_buf", "progress_buf"):
            assert torch.equal(out[k], ep[k][t]), f"step {t}: {k}"
        for k in ("obs", "rew", "potentials", "prev_potentials", "up_vec", "heading_vec"):
            torch.testing.assert_close(out[k], ep[k][t], rtol=1e-6, atol=1e-6, msg=f"step {t}: {k}")
        s["potentials"], s["prev_potentials"] = out["potentials"], out["prev_potentials"]
        s["reset_buf"], s["progress_buf"] = out["reset_buf"], out["progress_buf"]


@pytest.mark.skipif(not ref_loader.reference_available(), reason="reference tree only exists in the build container")
def test_episode_fixture_is_reproducible_from_the_reference():
    """Regenerating an episode from the verbatim classes gives the committed fixture."""
    from omniisaacgymenvs_b200.articulations import grid_env_origins
    from omniisaacgymenvs_b200.robots.model import build_model
    from oracle import ref_task_driver

    torch.set_num_threads(1)
    model = build_model("Ant")
    spawn = torch.from_numpy(grid_env_origins(96, 5.0))
    spawn[:, 2] += model.spawn_height
    ep = ref_task_driver.run_episode("Ant", 96, 5, 0, torch.from_numpy(model.dof_limits()), spawn)
    ref = _load("Ant")
    for k in ("reset_buf", "progress_buf", "reset_ids", "rand_pos", "after_reset_dof_pos", "efforts"):
        assert np.array_equal(ep[k], ref[k].numpy()), k
    for k in ("obs", "rew", "potentials"):
        assert np.allclose(ep[k], ref[k].numpy(), rtol=1e-6, atol=1e-6), k
