"""Observation / action noise oracle (oracle/dr_oracle.py) against the episode recorded from the VERBATIM reference
``Randomizer`` (utils/domain_randomization/randomize.py:212-305; tests/golden/episode_dr.npz written by
``python -m oracle.gen_golden dr``).  CPU only."""
import os

import numpy as np
import pytest
import torch

from oracle import dr_oracle as dro
from oracle import ref_loader

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "episode_dr.npz")
CASES = ("shadowhand", "scaling", "one_sided")


def load_case(name):
    z = np.load(GOLDEN)
    return {k.split("/", 1)[1]: z[k] for k in z.files if k.startswith(name + "/")}


def case_specs(name):
    from oracle.ref_dr_driver import CASES as BLOCKS, spec_of

    obs_block, act_block = BLOCKS[name]
    return (spec_of(obs_block, "on_reset"), spec_of(obs_block, "on_interval"),
            spec_of(act_block, "on_reset"), spec_of(act_block, "on_interval"))


def _rows(g, key, t):
    v = g[f"{key}_{t}"]
    return torch.from_numpy(v) if v.size or v.ndim == 2 else None


@pytest.mark.parametrize("name", CASES)
def test_oracle_replays_reference_episode_from_recorded_noise(name):
    """No dependence on torch's RNG stream: the noise rows recorded at generation time are fed back."""
    g = load_case(name)
    o_r, o_i, a_r, a_i = case_specs(name)
    steps, n, na = g["act_in"].shape
    an = dro.BufferNoise(n, na, a_r, a_i)
    on = dro.BufferNoise(n, g["obs_in"].shape[2], o_r, o_i)
    for t in range(steps):
        a, *_ = an.apply(torch.from_numpy(g["act_in"][t]).clone(), torch.from_numpy(g["pre_reset"][t]),
                         _rows(g, "act_noise_reset", t), _rows(g, "act_noise_interval", t))
        o, *_ = on.apply(torch.from_numpy(g["obs_in"][t]).clone(), torch.from_numpy(g["post_reset"][t]),
                         _rows(g, "obs_noise_reset", t), _rows(g, "obs_noise_interval", t))
        assert np.array_equal(a.numpy(), g["act_out"][t]) and np.array_equal(o.numpy(), g["obs_out"][t])
        assert np.array_equal(an.counter.numpy(), g["act_counter"][t]) and np.array_equal(on.counter.numpy(), g["obs_counter"][t])
        assert np.array_equal(an.correlated.numpy(), g["act_corr"][t]) and np.array_equal(on.correlated.numpy(), g["obs_corr"][t])


@pytest.mark.skipif(not ref_loader.reference_available(), reason="reference tree only exists in the build container")
@pytest.mark.parametrize("name", CASES)
def test_oracle_is_bit_identical_to_the_verbatim_randomizer(name):
    from oracle import ref_dr_driver

    blob = ref_dr_driver.run_case(name)  # asserts oracle == reference step by step with the same torch seed
    g = load_case(name)
    for k in ("act_out", "obs_out", "act_counter", "obs_counter"):
        assert np.array_equal(blob[k], g[k]), k


def test_counters_follow_the_reference_quirks():
    """counter[reset] = 0 happens BEFORE the += 1, and the interval test resets the counter to 0 (randomize.py:213-215,228-229)."""
    spec = dro.NoiseSpec("additive", "uniform", (0.0, 0.0), 3)
    bn = dro.BufferNoise(4, 2, None, spec)
    seen = []
    for t in range(7):
        rb = torch.tensor([0, 0, 0, 1 if t == 4 else 0])
        bn.apply(torch.zeros(4, 2), rb)
        seen.append(bn.counter.tolist())
    assert [s[0] for s in seen] == [1, 2, 0, 1, 2, 0, 1]
    assert [s[3] for s in seen] == [1, 2, 0, 1, 1, 2, 0]


@pytest.mark.parametrize("dist,params", [("gaussian", (0.5, 2.0)), ("uniform", (-0.3, 0.7)), ("loguniform", (0.5, 2.0))])
def test_philox_noise_distributions(dist, params):
    spec = dro.NoiseSpec("additive", dist, params)
    x = dro.philox_noise(spec, seed=7, counter=3, env_gids=np.arange(4096), dim=10, stream=2).astype(np.float64)
    assert x.shape == (4096, 10) and np.isfinite(x).all()
    if dist == "gaussian":
        assert abs(x.mean() - 0.5) < 0.05 and abs(x.std() - 2.0) < 0.05
        assert abs(((x - 0.5) ** 3).mean()) < 0.5  # symmetric
    elif dist == "uniform":
        assert x.min() >= -0.3 and x.max() < 0.7 and abs(x.mean() - 0.2) < 0.01
    else:
        assert x.min() >= 0.5 * (1 - 1e-6) and x.max() <= 2.0 * (1 + 1e-6) and abs(np.log(x).mean()) < 0.02
    # keyed by the GLOBAL env id: a shard draws the rows of the whole
    part = dro.philox_noise(spec, 7, 3, np.arange(1000, 1100), 10, 2)
    assert np.array_equal(part, x[1000:1100].astype(np.float32))
    # streams and counters are independent
    assert not np.array_equal(dro.philox_noise(spec, 7, 4, np.arange(100), 10, 2), x[:100].astype(np.float32))
    assert not np.array_equal(dro.philox_noise(spec, 7, 3, np.arange(100), 10, 3), x[:100].astype(np.float32))
