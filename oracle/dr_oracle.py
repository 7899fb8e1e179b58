"""TEST INFRASTRUCTURE ONLY -- CPU restatement (torch, fp32) of the reference's observation / action noise.

Follows /root/reference/omniisaacgymenvs/utils/domain_randomization/randomize.py:212-305 (``apply_observations_
randomization``, ``apply_actions_randomization``, ``_apply_uncorrelated_noise``, ``_apply_correlated_noise``) and the
two call sites in envs/vec_env_rlgames.py:59-60,70-71.  Pinned bit-exact against the VERBATIM reference ``Randomizer``
class (``oracle/ref_dr_driver.py`` -> ``tests/golden/episode_dr.npz``, ``tests/test_oracle_dr.py``).  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU legs may import this module; the product path never does.

Two ways to obtain the noise:
  * ``noise_reset`` / ``noise_interval`` left ``None``: drawn here with torch's global generator by the reference's own
    expressions in the reference's order (so the same ``torch.manual_seed`` reproduces the reference bit for bit);
  * given: ``[k, dim]`` rows for the k flagged envs in ascending env order -- the CUDA parity tests feed the SAME rows
    to the kernel (scattered to ``[N, dim]``), so kernel and oracle can be compared bit-exactly.
``philox_noise`` restates the in-kernel generator (csrc/domain_randomization.cu) for the sharding-invariant fast mode.
"""
from dataclasses import dataclass
from typing import Optional

import numpy as np
import torch

from oracle import philox

GAUSSIAN, UNIFORM, LOGUNIFORM = 0, 1, 2
ADDITIVE, SCALING = 0, 1
STREAM = {("actions", "on_reset"): 0, ("actions", "on_interval"): 1,
          ("observations", "on_reset"): 2, ("observations", "on_interval"): 3}


@dataclass(frozen=True)
class NoiseSpec:
    """One ``on_reset`` / ``on_interval`` block of the task YAML (cfg/task/ShadowHandOpenAI_FF.yaml:142-161)."""

    operation: str                    # "additive" | "scaling"
    distribution: str                 # "gaussian"/"normal" | "uniform" | "loguniform"/"log_uniform"
    distribution_parameters: tuple    # (mean, std) or (low, high)
    frequency_interval: int = 1       # on_interval only

    @property
    def dist_code(self) -> int:
        return {"gaussian": GAUSSIAN, "normal": GAUSSIAN, "uniform": UNIFORM, "loguniform": LOGUNIFORM,
                "log_uniform": LOGUNIFORM}[self.distribution]

    @property
    def op_code(self) -> int:
        return {"additive": ADDITIVE, "scaling": SCALING}[self.operation]


def draw(spec: NoiseSpec, rows: int, dim: int) -> torch.Tensor:
    """randomize.py:258-264 (and the identical :278-283), expressions kept verbatim."""
    distribution_parameters = spec.distribution_parameters
    if spec.dist_code == GAUSSIAN:
        return torch.normal(mean=distribution_parameters[0], std=distribution_parameters[1], size=(rows, dim))
    if spec.dist_code == UNIFORM:
        return (distribution_parameters[1] - distribution_parameters[0]) * torch.rand((rows, dim)) + distribution_parameters[0]
    return torch.exp((np.log(distribution_parameters[1]) - np.log(distribution_parameters[0])) * torch.rand((rows, dim))
                     + np.log(distribution_parameters[0]))


class BufferNoise:
    """State the reference keeps per randomised buffer (randomize.py:192-193 / :209-210)."""

    def __init__(self, num_envs: int, dim: int, on_reset: Optional[NoiseSpec], on_interval: Optional[NoiseSpec]):
        self.counter = torch.zeros(num_envs, dtype=torch.int32)
        self.correlated = torch.zeros(num_envs, dim, dtype=torch.float32)
        self.on_reset, self.on_interval = on_reset, on_interval

    def apply(self, buffer: torch.Tensor, reset_buf: torch.Tensor, noise_reset=None, noise_interval=None):
        """randomize.py:212-238 (observations) == :240-265 (actions).  In place on ``buffer``; returns
        (buffer, reset_ids, randomize_ids, noise_reset_used, noise_interval_used)."""
        env_ids = reset_buf.nonzero(as_tuple=False).squeeze(-1)
        self.counter[env_ids] = 0
        self.counter += 1
        used_r = used_i = None
        randomize_ids = torch.zeros(0, dtype=torch.int64)
        if self.on_reset is not None:  # _apply_correlated_noise, randomize.py:283-305
            if len(env_ids) > 0:
                used_r = draw(self.on_reset, len(env_ids), buffer.shape[1]) if noise_reset is None else noise_reset
                self.correlated[env_ids] = used_r
            if self.on_reset.op_code == ADDITIVE:
                buffer += self.correlated
            else:
                buffer *= self.correlated
        if self.on_interval is not None:  # _apply_uncorrelated_noise, randomize.py:267-281
            randomize_ids = (self.counter >= self.on_interval.frequency_interval).nonzero(as_tuple=False).squeeze(-1)
            self.counter[randomize_ids] = 0
            used_i = draw(self.on_interval, len(randomize_ids), buffer.shape[1]) if noise_interval is None else noise_interval
            if self.on_interval.op_code == ADDITIVE:
                buffer[randomize_ids] += used_i
            else:
                buffer[randomize_ids] *= used_i
        return buffer, env_ids, randomize_ids, used_r, used_i


def philox_noise(spec: NoiseSpec, seed: int, counter: int, env_gids, dim: int, stream: int) -> np.ndarray:
    """[len(env_gids), dim] float32 noise values the kernel generates for (seed, counter, global env id, stream).

    Element (env, col) has the GLOBAL flat index ``gid * dim + col``; one Philox4x32-10 call serves the group of 4
    consecutive flat indices: counter words (group_lo, group_hi, counter_lo, counter_hi), key (seed_lo,
    seed_hi ^ 0x44520000 ^ stream); the element takes component ``flat & 3``.  Uniform / log-uniform use the four
    24-bit uniforms; gaussian uses Box-Muller on the pairs (r0, r1) -> components 0, 1 and (r2, r3) -> components 2, 3
    with the radius uniform shifted to (0, 1]."""
    gid = np.asarray(env_gids, dtype=np.uint64)
    flat = (gid[:, None] * np.uint64(dim) + np.arange(dim, dtype=np.uint64)[None, :]).ravel()
    group, comp = flat >> np.uint64(2), (flat & np.uint64(3)).astype(np.int64)
    k = len(flat)
    c0 = (group & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    c1 = (group >> np.uint64(32)).astype(np.uint32)
    c2 = np.full(k, counter & 0xFFFFFFFF, np.uint32)
    c3 = np.full(k, (counter >> 32) & 0xFFFFFFFF, np.uint32)
    k0 = np.full(k, seed & 0xFFFFFFFF, np.uint32)
    k1 = np.full(k, ((seed >> 32) & 0xFFFFFFFF) ^ 0x44520000 ^ stream, np.uint32)
    p0, p1 = (np.float32(p) for p in spec.distribution_parameters)
    r = philox.philox4x32_10(c0, c1, c2, c3, k0, k1)
    if spec.dist_code == GAUSSIAN:
        vals = []
        for a, b in ((r[0], r[1]), (r[2], r[3])):
            u1 = ((a >> np.uint32(8)).astype(np.float32) + np.float32(1.0)) * np.float32(1.0 / 16777216.0)
            u2 = philox.u01(b)
            rad = np.sqrt(np.float32(-2.0) * np.log(u1)).astype(np.float32)
            ang = (np.float32(6.2831853071795864) * u2).astype(np.float32)
            vals += [rad * np.cos(ang).astype(np.float32), rad * np.sin(ang).astype(np.float32)]
        vals = [(p0 + p1 * v).astype(np.float32) for v in vals]
    elif spec.dist_code == UNIFORM:
        vals = [((p1 - p0) * philox.u01(x) + p0).astype(np.float32) for x in r]
    else:
        l0, l1 = np.float32(np.log(p0)), np.float32(np.log(p1))
        vals = [np.exp(((l1 - l0) * philox.u01(x) + l0).astype(np.float32)).astype(np.float32) for x in r]
    out = np.stack(vals, axis=1)[np.arange(k), comp]
    return out.reshape(len(gid), dim).astype(np.float32)
