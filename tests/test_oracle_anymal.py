"""CPU: the Anymal oracle (oracle/anymal_oracle.py) against the episode recorded from the VERBATIM reference
``AnymalTask`` / ``AnymalView`` (oracle/ref_anymal_driver.py -> tests/golden/episode_anymal.npz: post_reset, then
5 x [pre_physics_step, injected state, progress += 1, get_observations, calculate_metrics, is_done] through
tasks/anymal.py:137-293), and -- in the build container only -- against the reference run live.

The oracle calls the same torch ops in the same order as the reference, so the comparison is exact."""
import os

import numpy as np
import pytest
import torch

import helpers
from oracle import anymal_oracle as ao
from oracle import ref_loader


def load_episode():
    z = np.load(os.path.join(helpers.GOLDEN_DIR, "episode_anymal.npz"))
    return {k: torch.from_numpy(z[k]) for k in z.files}


def initial_state(ep):
    n = ep["actions"].shape[1]
    return {"pos": ep["spawn_pos"].clone(), "quat": ep["initial_root_rot"].clone(), "vel": torch.zeros(n, 6),
            "dof_pos": torch.zeros(n, 12), "dof_vel": torch.zeros(n, 12), "default_dof_pos": ep["default_dof_pos"],
            "current_targets": ep["default_dof_pos"].clone(), "commands": torch.zeros(n, 3),
            "actions": torch.zeros(n, 12), "last_actions": torch.zeros(n, 12), "last_dof_vel": torch.zeros(n, 12),
            "reset_buf": torch.ones(n, dtype=torch.int64), "progress_buf": torch.zeros(n, dtype=torch.int64),
            "initial_root_pos": ep["initial_root_pos"], "initial_root_rot": ep["initial_root_rot"],
            "dof_lo": ep["dof_lo"], "dof_hi": ep["dof_hi"]}


def replay(ep, spec=ao.ANYMAL):
    torch.set_num_threads(1)
    steps, n = ep["actions"].shape[:2]
    s = initial_state(ep)
    assert torch.equal(ep["initial_root_pos"], ep["spawn_pos"])
    assert torch.equal(ep["default_dof_pos"][0], spec.default_dof_pos())
    ao.reset_idx(spec, s, torch.arange(n), ep["post_reset_rand_vel"], ep["post_reset_rand_cmd"])
    assert torch.equal(s["dof_pos"], ep["post_reset_dof_pos"]) and torch.equal(s["dof_vel"], ep["post_reset_dof_vel"])
    assert torch.equal(s["commands"], ep["post_reset_commands"])
    for t in range(steps):
        ids = s["reset_buf"].nonzero(as_tuple=False).squeeze(-1)
        assert torch.equal(ids, ep["reset_mask"][t].nonzero(as_tuple=False).squeeze(-1)), f"step {t}: reset set"
        if len(ids):
            ao.reset_idx(spec, s, ids, ep["rand_vel"][t][ids], ep["rand_cmd"][t][ids])
        ao.integrate_targets(spec, s, ep["actions"][t])
        for k, g in (("dof_pos", "after_reset_dof_pos"), ("dof_vel", "after_reset_dof_vel"), ("pos", "after_reset_pos"),
                     ("quat", "after_reset_quat"), ("vel", "after_reset_vel"), ("commands", "commands"),
                     ("current_targets", "targets"), ("last_actions", "after_reset_last_actions"),
                     ("last_dof_vel", "after_reset_last_dof_vel")):
            assert torch.equal(s[k], ep[g][t]), f"step {t}: {k} after pre_physics_step"
        s.update(pos=ep["inj_pos"][t].clone(), quat=ep["inj_quat"][t].clone(), vel=ep["inj_vel"][t].clone(),
                 dof_pos=ep["inj_dof_pos"][t].clone(), dof_vel=ep["inj_dof_vel"][t].clone())
        if t == 2:
            s["progress_buf"][: n // 4] = spec.max_episode_length - 3
        out = ao.post_physics_step(spec, s)
        for k in ("reset_buf", "progress_buf"):
            assert torch.equal(out[k], ep[k][t]), f"step {t}: {k}"
        for k in ("obs", "rew", "last_actions", "last_dof_vel"):
            assert torch.equal(out[k], ep[k][t]), f"step {t}: {k}"


def test_spec_constants():
    spec = ao.ANYMAL
    assert spec.max_episode_length == 3000 and abs(spec.rew_scales()["lin_vel_xy"] - 1 / 60) < 1e-15
    assert ao.DOF_NAMES[:4] == ["LF_HAA", "LH_HAA", "RF_HAA", "RH_HAA"]  # the cosmetic reward's dof_pos[:, 0:4]


def test_oracle_reproduces_the_committed_reference_episode():
    ep = load_episode()
    assert (ep["rew"] == -1).any() and (ep["rew"] == 0).any() and (ep["rew"] > 0).any()
    assert (ep["progress_buf"] >= 2999).any() and (ep["obs"].abs() > 5.0).any()
    replay(ep)


@pytest.mark.skipif(not ref_loader.reference_available(), reason="reference tree only exists in the build container")
def test_oracle_matches_the_reference_run_live():
    from omniisaacgymenvs_b200.articulations import grid_env_origins
    from omniisaacgymenvs_b200.robots.model import build_model
    from oracle import ref_anymal_driver

    model = build_model("Anymal")
    spawn = torch.from_numpy(grid_env_origins(40, 4.0))
    spawn[:, 2] += model.spawn_height
    ep = ref_anymal_driver.run_episode(40, steps=4, seed=5, spawn_pos=spawn, dof_limits=torch.from_numpy(model.dof_limits()))
    replay({k: torch.from_numpy(v) for k, v in ep.items()})


@pytest.mark.skipif(not ref_loader.reference_available(), reason="reference tree only exists in the build container")
def test_committed_fixture_is_what_the_reference_produces():
    from omniisaacgymenvs_b200.robots.model import build_model
    from oracle import ref_anymal_driver

    ep = load_episode()
    model = build_model("Anymal")
    live = ref_anymal_driver.run_episode(96, steps=5, seed=0, spawn_pos=ep["spawn_pos"].clone(),
                                         dof_limits=torch.from_numpy(model.dof_limits()))
    for k, v in live.items():
        assert np.array_equal(ep[k].numpy(), v), k
