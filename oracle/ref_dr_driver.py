"""TEST INFRASTRUCTURE ONLY -- drives the VERBATIM reference ``Randomizer`` (observation / action noise) on the CPU.

``utils/domain_randomization/randomize.py`` is executed unmodified from /root/reference (never copied) after
``sys.modules`` stubs for ``omni.replicator.core`` / ``omni.replicator.isaac`` / ``omni.isaac.core.prims`` are
installed (the replicator trigger and orchestrator become no-ops; only the ``observations`` and ``actions`` keys of
``randomization_params`` are used, which touch nothing else).  ``Randomizer.__init__`` (:40-57),
``set_up_domain_randomization`` (:126-166), ``apply_actions_randomization`` (:240-265) and
``apply_observations_randomization`` (:212-238) then run exactly as written, in the order of
``VecEnvRLGames.step`` (envs/vec_env_rlgames.py:59-60,70-71).  Container-only (needs /root/reference);
``oracle/gen_golden.py`` uses it to write ``tests/golden/episode_dr.npz``.
"""
import contextlib
from typing import Dict

import numpy as np
import torch

from oracle import dr_oracle as dro
from oracle import ref_loader

# (name, observations block, actions block): every distribution x operation, interval > 1, one-sided blocks
CASES = {
    "shadowhand": (  # cfg/task/ShadowHandOpenAI_FF.yaml:142-161
        {"on_reset": dict(operation="additive", distribution="gaussian", distribution_parameters=[0, .0001]),
         "on_interval": dict(frequency_interval=1, operation="additive", distribution="gaussian", distribution_parameters=[0, .002])},
        {"on_reset": dict(operation="additive", distribution="gaussian", distribution_parameters=[0, 0.015]),
         "on_interval": dict(frequency_interval=1, operation="additive", distribution="gaussian", distribution_parameters=[0., 0.05])}),
    "scaling": (
        {"on_reset": dict(operation="scaling", distribution="uniform", distribution_parameters=[0.9, 1.1]),
         "on_interval": dict(frequency_interval=3, operation="scaling", distribution="loguniform", distribution_parameters=[0.8, 1.25])},
        {"on_reset": dict(operation="scaling", distribution="log_uniform", distribution_parameters=[0.5, 2.0]),
         "on_interval": dict(frequency_interval=2, operation="additive", distribution="uniform", distribution_parameters=[-0.1, 0.1])}),
    "one_sided": (
        {"on_interval": dict(frequency_interval=2, operation="additive", distribution="normal", distribution_parameters=[0.1, 0.3])},
        {"on_reset": dict(operation="additive", distribution="uniform", distribution_parameters=[-0.05, 0.05])}),
}


def _install_dr_stubs() -> None:
    ref_loader.install_stubs()
    trigger = type("trigger", (), {"on_rl_frame": staticmethod(lambda **k: contextlib.nullcontext())})
    ref_loader._mod("omni.replicator")
    ref_loader._mod("omni.replicator.core", set_global_seed=lambda s: None,
                    orchestrator=type("orchestrator", (), {"run": staticmethod(lambda: None)}))
    ref_loader._mod("omni.replicator.isaac", trigger=trigger)
    ref_loader._mod("omni.isaac.core.prims", RigidPrimView=ref_loader._Dummy)


def load_randomizer_class():
    _install_dr_stubs()
    mod = ref_loader._load("omniisaacgymenvs.utils.domain_randomization.randomize",
                           "omniisaacgymenvs/utils/domain_randomization/randomize.py")
    return mod.Randomizer


class _SimConfig:
    def __init__(self, n, obs_block, act_block, seed):
        self.task_config = {"env": {"numEnvs": n},
                            "domain_randomization": {"randomize": True, "min_frequency": 1,
                                                     "randomization_params": {"observations": obs_block, "actions": act_block}}}
        self.config = {"sim_device": "cpu", "seed": seed}


class _Task:
    def __init__(self, num_obs, num_actions):
        self.num_observations, self.num_actions = num_obs, num_actions
        self.randomize_observations = self.randomize_actions = False


def spec_of(block, key):
    if block is None or key not in block:
        return None
    b = block[key]
    return dro.NoiseSpec(b["operation"], b["distribution"], tuple(float(x) for x in b["distribution_parameters"]),
                         int(b.get("frequency_interval", 1)))


def run_case(name: str, n: int = 33, num_obs: int = 60, num_actions: int = 8, steps: int = 7, seed: int = 0) -> Dict[str, np.ndarray]:
    """Reference episode + the oracle's replay of it (asserted bit-identical here, at generation time)."""
    Randomizer = load_randomizer_class()
    obs_block, act_block = CASES[name]
    gen = torch.Generator().manual_seed(1000 + seed)
    acts_in = [torch.rand(n, num_actions, generator=gen) * 2 - 1 for _ in range(steps)]
    obs_in = [torch.randn(n, num_obs, generator=gen) for _ in range(steps)]
    # reset_buf seen by the actions hook (before pre_physics_step) and by the observations hook (after is_done):
    # all envs flagged on the first step (RLTask.reset), none on the third, a few otherwise
    pre = [(torch.rand(n, generator=gen) < 0.2).long() for _ in range(steps)]
    post = [(torch.rand(n, generator=gen) < 0.2).long() for _ in range(steps)]
    pre[0][:] = 1
    pre[2][:] = 0
    post[2][:] = 0

    task = _Task(num_obs, num_actions)
    r = Randomizer(_SimConfig(n, obs_block, act_block, seed))
    r.set_up_domain_randomization(task)
    assert task.randomize_observations and task.randomize_actions
    torch.manual_seed(seed)
    out: Dict[str, list] = {k: [] for k in ("act_out", "obs_out", "act_counter", "obs_counter", "act_corr", "obs_corr")}
    for t in range(steps):
        a = r.apply_actions_randomization(actions=acts_in[t].clone(), reset_buf=pre[t])
        o = r.apply_observations_randomization(observations=obs_in[t].clone(), reset_buf=post[t])
        out["act_out"].append(a.clone()); out["obs_out"].append(o.clone())
        out["act_counter"].append(r._actions_counter_buffer.clone()); out["obs_counter"].append(r._observations_counter_buffer.clone())
        out["act_corr"].append(r._actions_correlated_noise.clone()); out["obs_corr"].append(r._observations_correlated_noise.clone())

    # oracle replay with the same seed: must be bit-identical; it also records the noise rows it drew
    an = dro.BufferNoise(n, num_actions, spec_of(act_block, "on_reset"), spec_of(act_block, "on_interval"))
    on = dro.BufferNoise(n, num_obs, spec_of(obs_block, "on_reset"), spec_of(obs_block, "on_interval"))
    torch.manual_seed(seed)
    blob: Dict[str, np.ndarray] = {}
    for t in range(steps):
        a, _, _, ar, ai = an.apply(acts_in[t].clone(), pre[t])
        o, _, _, orr, oi = on.apply(obs_in[t].clone(), post[t])
        for got, ref in ((a, out["act_out"][t]), (o, out["obs_out"][t]), (an.counter, out["act_counter"][t]),
                         (on.counter, out["obs_counter"][t]), (an.correlated, out["act_corr"][t]),
                         (on.correlated, out["obs_corr"][t])):
            assert torch.equal(got, ref), f"oracle deviates from the reference Randomizer ({name}, step {t})"
        for key, v in (("act_noise_reset", ar), ("act_noise_interval", ai), ("obs_noise_reset", orr), ("obs_noise_interval", oi)):
            blob[f"{key}_{t}"] = (v if v is not None else torch.zeros(0)).numpy()
    blob.update({k: torch.stack(v).numpy() for k, v in out.items()})
    blob.update(act_in=torch.stack(acts_in).numpy(), obs_in=torch.stack(obs_in).numpy(),
                pre_reset=torch.stack(pre).numpy(), post_reset=torch.stack(post).numpy(),
                seed=np.int64(seed), case=np.array(name))
    return blob


def run_all() -> Dict[str, np.ndarray]:
    blob = {}
    for name in CASES:
        for k, v in run_case(name).items():
            blob[f"{name}/{k}"] = v
    return blob
