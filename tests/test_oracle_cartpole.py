"""CPU: the Cartpole oracle (oracle/cartpole_oracle.py) against the episode recorded from the VERBATIM
reference ``CartpoleTask`` (oracle/ref_cartpole_driver.py -> tests/golden/episode_cartpole.npz: post_reset, then
6 x [pre_physics_step, injected joint state, progress += 1, get_observations, calculate_metrics, is_done]
through tasks/cartpole.py:80-162), and -- in the build container only -- against the reference run live."""
import os

import numpy as np
import pytest
import torch

import helpers
from oracle import cartpole_oracle as co
from oracle import ref_loader


def _load():
    z = np.load(os.path.join(helpers.GOLDEN_DIR, "episode_cartpole.npz"))
    return {k: torch.from_numpy(z[k]) for k in z.files}


def _replay(ep):
    """Replays the episode with the oracle; every quantity must equal the recording BIT FOR BIT (all of it is
    elementwise fp32 arithmetic with no reductions, so there is no summation-order freedom)."""
    spec = co.CARTPOLE
    steps, n = ep["actions"].shape[:2]
    q, qd = torch.zeros(n, 2), torch.zeros(n, 2)
    reset_buf = torch.ones(n, dtype=torch.int64)
    progress = torch.zeros(n, dtype=torch.int64)
    # post_reset: reset_idx(all) (cartpole.py:132-137)
    q[:], qd[:] = co.reset_values(ep["post_reset_rand"])
    reset_buf[:], progress[:] = 0, 0
    assert torch.equal(q, ep["post_reset_dof_pos"]) and torch.equal(qd, ep["post_reset_dof_vel"])
    for t in range(steps):
        ids = reset_buf.nonzero(as_tuple=False).squeeze(-1)
        assert torch.equal(ids, ep["reset_mask"][t].nonzero(as_tuple=False).squeeze(-1)), f"step {t}: reset set"
        if len(ids):
            q[ids], qd[ids] = co.reset_values(ep["rand"][t][ids])
            reset_buf[ids], progress[ids] = 0, 0
        assert torch.equal(q, ep["after_reset_dof_pos"][t]) and torch.equal(qd, ep["after_reset_dof_vel"][t])
        assert torch.equal(co.efforts(spec, ep["actions"][t]), ep["efforts"][t])
        q, qd = ep["inj_dof_pos"][t].clone(), ep["inj_dof_vel"][t].clone()
        if t == 2:
            progress[: n // 4] = 498
        obs, rew, reset_buf, progress = co.post_physics_step(spec, q, qd, progress)
        assert torch.equal(obs, ep["obs"][t]), f"step {t}: obs"
        assert torch.equal(rew, ep["rew"][t]), f"step {t}: rew"
        assert torch.equal(reset_buf, ep["reset_buf"][t]), f"step {t}: reset_buf"
        assert torch.equal(progress, ep["progress_buf"][t]), f"step {t}: progress_buf"


def test_oracle_reproduces_the_committed_reference_episode():
    ep = _load()
    # the fixture exercises every branch
    assert (ep["rew"] == -2.0).any() and (ep["rew"] > 0).any()
    assert (ep["progress_buf"] >= 500).any() and (ep["obs"].abs() > 5.0).any()
    _replay(ep)


@pytest.mark.skipif(not ref_loader.reference_available(), reason="reference tree only exists in the build container")
def test_oracle_matches_the_reference_run_live():
    from oracle import ref_cartpole_driver

    ep = {k: torch.from_numpy(v) for k, v in ref_cartpole_driver.run_episode(64, steps=4, seed=7).items()}
    _replay(ep)


@pytest.mark.skipif(not ref_loader.reference_available(), reason="reference tree only exists in the build container")
def test_committed_fixture_is_what_the_reference_produces():
    from oracle import ref_cartpole_driver

    ep, live = _load(), ref_cartpole_driver.run_episode(96, steps=6, seed=0)
    for k, v in live.items():
        assert np.array_equal(ep[k].numpy(), v), k


def test_boundary_comparisons_are_strict_and_in_fp32():
    spec = co.CARTPOLE
    q, qd = co.synthetic_state(16, 0)
    obs, rew, reset, _ = co.post_physics_step(spec, q, qd, torch.zeros(16, dtype=torch.int64))
    # |cart| == 3.0 exactly -> no reset; one ulp above -> reset
    assert reset[0] == 0 or abs(float(q[0, 1])) > np.pi / 2
    assert reset[2] == 1 and rew[2] == -2.0
    # |pole| == fp32(pi/2) compares equal to the fp32-rounded threshold -> no reset from the angle
    only_angle = lambda i: abs(float(q[i, 0])) <= 3.0
    if only_angle(3):
        assert reset[3] == 0
    assert reset[4] == 1 and reset[5] == 1
