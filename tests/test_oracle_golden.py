"""CPU: the oracle restatement against (a) the committed golden vectors, which were produced by
the VERBATIM reference functions (oracle/gen_golden.py), and (b) the verbatim reference itself
when /root/reference is present (build container only)."""
import pytest
import torch

import helpers
from oracle import locomotion_oracle as lo
from oracle import ref_loader


@pytest.mark.parametrize("name", helpers.golden_names())
def test_oracle_reproduces_golden(name):
    torch.set_num_threads(1)
    s, ref = helpers.load_golden(name)
    spec = lo.SPECS["Humanoid" if name.startswith("humanoid") else "Ant"]
    got = lo.post_physics_step(spec, {k: v.clone() for k, v in s.items()})
    assert torch.equal(got["reset_buf"], ref["reset_buf"])
    assert torch.equal(got["progress_buf"], ref["progress_buf"])
    # torch's CPU kernels pick vectorised or scalar code by tensor size / thread count, so float
    # results are only guaranteed to tolerance across machines (bit-exact on the generating box)
    for k in ("obs", "rew", "potentials", "prev_potentials", "up_vec", "heading_vec"):
        torch.testing.assert_close(got[k], ref[k], rtol=1e-6, atol=1e-6, msg=k)


def test_golden_set_is_complete():
    names = helpers.golden_names()
    for robot in ("humanoid", "ant"):
        assert f"{robot}_edge" in names
        for seed in range(4):
            for n in (1, 33, 256):
                assert f"{robot}_s{seed}_n{n}" in names


@pytest.mark.skipif(not ref_loader.reference_available(), reason="reference tree only exists in the build container")
@pytest.mark.parametrize("robot", ["Humanoid", "Ant"])
@pytest.mark.parametrize("n", [1, 33, 4096])
def test_oracle_equals_verbatim_reference(robot, n):
    from oracle.gen_golden import run_reference

    spec = lo.SPECS[robot]
    s = lo.synthetic_state(spec, n, 7)
    ref = run_reference(spec, {k: v.clone() for k, v in s.items()})
    got = lo.post_physics_step(spec, {k: v.clone() for k, v in s.items()})
    for k in ref:
        assert torch.equal(got[k], ref[k]), k


def test_edge_rows_hit_every_branch():
    for robot in ("humanoid", "ant"):
        s, ref = helpers.load_golden(f"{robot}_edge")
        spec = lo.SPECS[robot.capitalize()]
        # termination exactly at the threshold is NOT a fall (strict <), one ulp below is
        assert int(ref["reset_buf"][0]) == 0 and int(ref["reset_buf"][2]) == 1
        assert float(ref["rew"][2]) == spec.death_cost
        # progress 998 -> 999 after the increment -> reset; 997 -> 998 -> no reset
        assert [int(v) for v in ref["reset_buf"][4:8]] == [0, 0, 1, 1]
        assert int(ref["reset_buf"][32]) == 1  # incoming reset flag is preserved
        assert torch.isfinite(ref["obs"]).all() and torch.isfinite(ref["rew"]).all()
