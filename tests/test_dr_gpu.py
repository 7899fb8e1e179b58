"""Observation / action noise kernel (csrc/domain_randomization.cu, C-ABI ``oige_dr_apply_f32``) against
oracle/dr_oracle.py and the episode recorded from the verbatim reference ``Randomizer``
(utils/domain_randomization/randomize.py:212-305; tests/golden/episode_dr.npz)."""
import ctypes as C
import math

import numpy as np
import pytest
import torch

from omniisaacgymenvs_b200 import _lib
from oracle import dr_oracle as dro
from test_oracle_dr import CASES, case_specs, load_case

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def c_spec(s):
    if s is None:
        return _lib.OigeNoiseSpec(0, 0, 0, 0, 0.0, 0.0)
    return _lib.OigeNoiseSpec(1, s.op_code, s.dist_code, s.frequency_interval, *s.distribution_parameters)


class KernelNoise:
    """Device-side twin of dro.BufferNoise."""

    def __init__(self, n, dim, on_reset, on_interval, stream_id, seed=0, env_id_offset=0):
        self.n, self.dim = n, dim
        self.r, self.i = c_spec(on_reset), c_spec(on_interval)
        self.counter = torch.zeros(n, dtype=torch.int32, device=DEV)
        self.correlated = torch.zeros(n, dim, device=DEV)
        self.state = _lib.OigeNoiseState(self.counter.data_ptr(), self.correlated.data_ptr())
        self.stream_id, self.seed, self.off = stream_id, seed, env_id_offset

    def apply(self, lib, buf, reset_buf, noise_reset=None, noise_interval=None, counter=0, clip_in=math.inf,
              clip_out=math.inf, out=None, counter_dev=None):
        out = buf if out is None else out
        rc = lib.oige_dr_apply_f32(C.byref(self.r), C.byref(self.i), C.byref(self.state), buf.data_ptr(), out.data_ptr(),
                                   reset_buf.data_ptr(), buf.shape[0], self.dim, clip_in, clip_out,
                                   None if noise_reset is None else noise_reset.data_ptr(),
                                   None if noise_interval is None else noise_interval.data_ptr(), self.seed, counter,
                                   None if counter_dev is None else counter_dev.data_ptr(), self.stream_id, self.off,
                                   _lib.current_stream_ptr(torch.device(DEV)))
        _lib.check(rc, "oige_dr_apply_f32")
        return out


def scatter(rows, ids, n, dim):
    """[k, dim] noise rows of the flagged envs -> [N, dim] with NaN elsewhere (the kernel must not read those)."""
    full = torch.full((n, dim), float("nan"))
    if rows is not None and len(ids):
        full[ids] = rows
    return full.to(DEV)


@pytest.mark.parametrize("name", CASES)
def test_kernel_reproduces_reference_episode_bit_exact(cuda_lib, name):
    g = load_case(name)
    o_r, o_i, a_r, a_i = case_specs(name)
    steps, n, na = g["act_in"].shape
    no = g["obs_in"].shape[2]
    ka = KernelNoise(n, na, a_r, a_i, _lib.DR_STREAM_ACTIONS)
    ko = KernelNoise(n, no, o_r, o_i, _lib.DR_STREAM_OBSERVATIONS)
    oa, oo = dro.BufferNoise(n, na, a_r, a_i), dro.BufferNoise(n, no, o_r, o_i)
    for t in range(steps):
        for kern, orc, key, rkey in ((ka, oa, "act", "pre_reset"), (ko, oo, "obs", "post_reset")):
            rb = torch.from_numpy(g[rkey][t])
            rows_r = g[f"{key}_noise_reset_{t}"]
            rows_i = g[f"{key}_noise_interval_{t}"]
            rows_r = torch.from_numpy(rows_r) if rows_r.ndim == 2 else None
            rows_i = torch.from_numpy(rows_i) if rows_i.ndim == 2 else None
            # the oracle tells which envs the rows belong to
            _, ids_r, ids_i, _, _ = orc.apply(torch.from_numpy(g[f"{key}_in"][t]).clone(), rb, rows_r, rows_i)
            buf = torch.from_numpy(g[f"{key}_in"][t]).to(DEV)
            kern.apply(cuda_lib, buf, rb.to(DEV), scatter(rows_r, ids_r, n, kern.dim), scatter(rows_i, ids_i, n, kern.dim))
            assert np.array_equal(buf.cpu().numpy(), g[f"{key}_out"][t]), (key, t)
            assert np.array_equal(kern.counter.cpu().numpy(), g[f"{key}_counter"][t]), (key, t)
            assert np.array_equal(kern.correlated.cpu().numpy(), g[f"{key}_corr"][t]), (key, t)


@pytest.mark.parametrize("dist,params,op", [("gaussian", (0.1, 0.5), "additive"), ("uniform", (0.9, 1.1), "scaling"),
                                            ("loguniform", (0.5, 2.0), "scaling"), ("normal", (0.0, 0.02), "additive")])
@pytest.mark.parametrize("n,dim,off", [(1, 8, 100_000), (77, 87, 100_000), (4096, 60, 100_000), (1000, 21, 100_000),
                                       (333, 87, 3), (130, 21, 1)])
def test_philox_mode_matches_numpy_restatement(cuda_lib, dist, params, op, n, dim, off):
    """In-kernel Philox noise == oracle/dr_oracle.philox_noise fed through the oracle's apply.  Offsets 3 and 1 with an
    odd row length put the shard on a global flat index that is not a multiple of 4: the element-wise path."""
    r_spec = dro.NoiseSpec(op, dist, params)
    i_spec = dro.NoiseSpec(op, dist, params, 2)
    seed = 0x1234_5678_9ABC
    kern = KernelNoise(n, dim, r_spec, i_spec, _lib.DR_STREAM_OBSERVATIONS, seed, off)
    orc = dro.BufferNoise(n, dim, r_spec, i_spec)
    gen = torch.Generator().manual_seed(n * 131 + dim)
    gids = np.arange(off, off + n)
    for t in range(5):
        ctr = (1 << 33) + 17 + t
        buf = torch.randn(n, dim, generator=gen)
        rb = (torch.rand(n, generator=gen) < (1.0 if t == 0 else 0.3)).long()
        full_r = torch.from_numpy(dro.philox_noise(r_spec, seed, ctr, gids, dim, _lib.DR_STREAM_OBSERVATIONS))
        full_i = torch.from_numpy(dro.philox_noise(i_spec, seed, ctr, gids, dim, _lib.DR_STREAM_OBSERVATIONS + 1))
        ids_r = rb.nonzero().squeeze(-1)
        cnt = orc.counter.clone(); cnt[ids_r] = 0; cnt += 1
        ids_i = (cnt >= 2).nonzero().squeeze(-1)
        want, *_ = orc.apply(buf.clone(), rb, full_r[ids_r], full_i[ids_i])
        got = kern.apply(cuda_lib, buf.to(DEV), rb.to(DEV), counter=ctr).cpu()
        # device logf / sincosf / expf vs numpy: a few ulp of the noise value
        scale = max(abs(params[0]), abs(params[1]), 1.0)
        assert torch.allclose(got, want, rtol=2e-6, atol=4e-6 * scale), float((got - want).abs().max())
        assert torch.equal(kern.counter.cpu(), orc.counter)
        assert torch.allclose(kern.correlated.cpu(), orc.correlated, rtol=2e-6, atol=4e-6 * scale)


def test_clamps_sit_before_and_after_the_noise(cuda_lib):
    """clip_in = the wrapper's action clamp (vec_env_rlgames.py:57), clip_out = its observation clamp (:42)."""
    n, dim = 64, 5
    spec = dro.NoiseSpec("additive", "uniform", (0.5, 0.5))  # constant +0.5
    kern = KernelNoise(n, dim, None, spec, _lib.DR_STREAM_ACTIONS)
    buf = torch.linspace(-3, 3, n * dim).reshape(n, dim).to(DEV)
    rb = torch.zeros(n, dtype=torch.int64, device=DEV)
    out = torch.empty_like(buf)
    kern.apply(cuda_lib, buf, rb, clip_in=1.0, out=out)
    assert torch.equal(out, buf.clamp(-1, 1) + 0.5) and not torch.equal(out, buf)
    kern.apply(cuda_lib, buf, rb, clip_out=1.0, out=out)
    assert torch.equal(out, (buf + 0.5).clamp(-1, 1))


def test_unaligned_views_take_the_elementwise_path(cuda_lib):
    """A buffer view whose base is only 4-byte aligned gives the same result as the aligned tensor."""
    n, dim = 200, 87
    spec = dro.NoiseSpec("additive", "gaussian", (0.0, 0.1), 1)
    a = KernelNoise(n, dim, spec, spec, _lib.DR_STREAM_OBSERVATIONS, seed=5)
    b = KernelNoise(n, dim, spec, spec, _lib.DR_STREAM_OBSERVATIONS, seed=5)
    x = torch.randn(n, dim, device=DEV)
    pad = torch.zeros(n * dim + 1, device=DEV)
    y = pad[1:].view(n, dim)
    y.copy_(x)
    assert y.data_ptr() % 16 != 0
    rb = (torch.rand(n, device=DEV) < 0.5).long()
    a.apply(cuda_lib, x, rb, counter=4)
    b.apply(cuda_lib, y, rb, counter=4)
    assert torch.equal(x, y) and torch.equal(a.correlated, b.correlated)


def test_noise_is_keyed_by_global_env_id(cuda_lib):
    """A shard [a, b) with env_id_offset = a produces the rows of the un-sharded run, bit for bit."""
    n, dim, a, b = 1000, 60, 352, 704
    r_spec = dro.NoiseSpec("additive", "gaussian", (0.0, 0.1))
    i_spec = dro.NoiseSpec("scaling", "loguniform", (0.8, 1.25), 1)
    whole = KernelNoise(n, dim, r_spec, i_spec, _lib.DR_STREAM_OBSERVATIONS, seed=9)
    part = KernelNoise(b - a, dim, r_spec, i_spec, _lib.DR_STREAM_OBSERVATIONS, seed=9, env_id_offset=a)
    gen = torch.Generator().manual_seed(3)
    for t in range(4):
        buf = torch.randn(n, dim, generator=gen).to(DEV)
        rb = (torch.rand(n, generator=gen) < (1.0 if t == 0 else 0.2)).long().to(DEV)
        pb = buf[a:b].clone()
        whole.apply(cuda_lib, buf, rb, counter=t + 1)
        part.apply(cuda_lib, pb, rb[a:b].contiguous(), counter=t + 1)
        assert torch.equal(pb, buf[a:b]) and torch.equal(part.correlated, whole.correlated[a:b])


def test_argument_errors(cuda_lib):
    n, dim = 8, 4
    buf = torch.zeros(n, dim, device=DEV); rb = torch.zeros(n, dtype=torch.int64, device=DEV)
    kern = KernelNoise(n, dim, dro.NoiseSpec("additive", "uniform", (0.0, 1.0)), None, _lib.DR_STREAM_ACTIONS)
    kern.r.distribution = 7
    with pytest.raises(RuntimeError, match="distribution"):
        kern.apply(cuda_lib, buf, rb)
    kern.r.distribution = _lib.DR_LOGUNIFORM; kern.r.p0 = -1.0
    with pytest.raises(RuntimeError, match="loguniform"):
        kern.apply(cuda_lib, buf, rb)
    kern.r = c_spec(dro.NoiseSpec("additive", "uniform", (0.0, 1.0)))
    kern.state.correlated = None
    with pytest.raises(RuntimeError, match="correlated"):
        kern.apply(cuda_lib, buf, rb)
    kern.state.correlated = kern.correlated.data_ptr()
    kern.stream_id = 5
    with pytest.raises(RuntimeError, match="stream_id"):
        kern.apply(cuda_lib, buf, rb)


# ---- through the env wrapper ---------------------------------------------------------------------------------
DR_BLOCK = {"randomize": True, "min_frequency": 1, "randomization_params": {
    "observations": {"on_reset": {"operation": "additive", "distribution": "gaussian", "distribution_parameters": [0, 0.01]},
                     "on_interval": {"frequency_interval": 1, "operation": "additive", "distribution": "gaussian",
                                     "distribution_parameters": [0, 0.05]}},
    "actions": {"on_reset": {"operation": "additive", "distribution": "gaussian", "distribution_parameters": [0, 0.015]},
                "on_interval": {"frequency_interval": 1, "operation": "additive", "distribution": "uniform",
                                "distribution_parameters": [-0.05, 0.05]}}}}


def _env(task, n, dr=True, **kw):
    from omniisaacgymenvs_b200.utils.task_util import make_env

    ov = {"task.domain_randomization": DR_BLOCK} if dr else {}
    env, t = make_env(task, num_envs=n, overrides=ov, **kw)
    env.reset()
    return env, t


@pytest.mark.parametrize("task_name", ["Ant", "Humanoid", "Anymal", "Cartpole"])
def test_env_step_with_noise_eager_equals_graph(cuda_lib, task_name):
    n = 256
    env_e, task_e = _env(task_name, n)
    env_g, task_g = _env(task_name, n)
    assert task_e.randomize_actions and task_e.randomize_observations
    env_g.enable_cuda_graph()
    gen = torch.Generator(device=DEV).manual_seed(5)
    for t in range(12):
        a = torch.rand(n, task_e.num_actions, device=DEV, generator=gen) * 2 - 1
        oe, re_, de, _ = env_e.step(a)
        og, rg, dg, _ = env_g.step(a)
        assert torch.isfinite(oe["obs"]).all()
        assert torch.equal(oe["obs"], og["obs"]) and torch.equal(re_, rg) and torch.equal(de, dg), t
    assert float(oe["obs"].abs().max()) <= float(task_e.clip_obs)


def test_rebinding_a_buffer_in_graph_mode_drops_the_graph_and_keeps_the_noise_moving(cuda_lib):
    """A caller re-assigns `task.obs_buf` while the step is captured (the reference's reset() re-creates buffers): the
    graphs are dropped with ONE warning, the Randomizer goes back to the host-side step counter, and the eager steps
    that follow draw the same fresh noise an always-eager env draws -- not the same draw every step from the device
    counter nobody advances any more."""
    n = 256
    env_e, task_e = _env("Ant", n)
    env_g, task_g = _env("Ant", n)
    env_g.enable_cuda_graph()
    gen = torch.Generator(device=DEV).manual_seed(9)

    def both(t):
        a = torch.rand(n, task_e.num_actions, device=DEV, generator=gen) * 2 - 1
        oe, re_, de, _ = env_e.step(a)
        og, rg, dg, _ = env_g.step(a)
        assert torch.equal(oe["obs"], og["obs"]) and torch.equal(re_, rg) and torch.equal(de, dg), t
        return og["obs"].clone()

    for t in range(3):
        both(t)
    assert env_g._graph is not None
    for task in (task_e, task_g):
        task.obs_buf = task.obs_buf.clone()
    task_e.get_observations()
    with pytest.warns(RuntimeWarning, match="captured graphs were dropped"):
        task_g.get_observations()  # the first hook after the re-assignment notices it
    assert env_g._graph is None and task_g._dr_randomizer._counter_dev is None
    seen = [both(t) for t in range(3, 8)]
    assert all(not torch.equal(seen[i], seen[i + 1]) for i in range(len(seen) - 1))
    import warnings as _w
    with _w.catch_warnings():
        _w.simplefilter("error")  # warned once, not on every later hook
        task_g.obs_buf = task_g.obs_buf.clone()
        task_g.get_observations()


def test_env_noise_changes_observations_by_the_configured_amount(cuda_lib):
    """Only observation noise active: the physics is identical to the noise-free env, the returned observations
    differ from it by correlated + uncorrelated gaussians of the configured sigmas."""
    from omniisaacgymenvs_b200.utils.task_util import make_env

    n = 4096
    block = {"randomize": True, "randomization_params": {"observations": DR_BLOCK["randomization_params"]["observations"]}}
    env_n, task_n = make_env("Humanoid", num_envs=n, overrides={"task.domain_randomization": block}); env_n.reset()
    env_c, task_c = make_env("Humanoid", num_envs=n); env_c.reset()
    assert task_n.randomize_observations and not task_n.randomize_actions and not task_c.randomize_observations
    gen = torch.Generator(device=DEV).manual_seed(11)
    ever_done = torch.zeros(n, dtype=torch.bool, device=DEV)
    for _ in range(25):
        a = torch.rand(n, 21, device=DEV, generator=gen) * 2 - 1
        on, _, dn, _ = env_n.step(a)
        oc, _, dc, _ = env_c.step(a)
        assert torch.equal(dn, dc)
        ever_done |= dn.bool()
    assert torch.equal(task_n.get_robot().dof_pos, task_c.get_robot().dof_pos)
    corr = task_n._dr_randomizer._observations_correlated_noise
    # on_reset noise is drawn when the hook sees reset_buf != 0 (after is_done), i.e. never before an env's first
    # termination (randomize.py:213,291-292); afterwards it stays until the next one
    assert int(ever_done.sum()) > 200 and int((~ever_done).sum()) > 200, float(ever_done.float().mean())
    assert float(corr[~ever_done].abs().max()) == 0.0
    assert abs(float(corr[ever_done].std()) - 0.01) < 0.0005
    inside = (oc["obs"].abs() < 4.5) & (on["obs"].abs() < 4.9)  # away from the +-5 observation clamp
    resid = (on["obs"] - oc["obs"] - corr)[inside]
    assert abs(float(resid.std()) - 0.05) < 0.001 and abs(float(resid.mean())) < 0.0005


def test_unsupported_randomisation_keys_are_rejected(cuda_lib):
    from omniisaacgymenvs_b200.utils.task_util import make_env

    block = {"randomize": True, "randomization_params": {"simulation": {"gravity": {"on_interval": {
        "frequency_interval": 720, "operation": "additive", "distribution": "gaussian",
        "distribution_parameters": [[0, 0, 0.0], [0, 0, 0.4]]}}}}}
    with pytest.raises(RuntimeError, match="not part of this build"):
        make_env("Ant", num_envs=64, overrides={"task.domain_randomization": block})
    bad = {"randomize": True, "randomization_params": {"observations": {"on_reset": {"operation": "additive"}}}}
    with pytest.raises(ValueError, match="on_reset randomization parameters"):
        make_env("Ant", num_envs=64, overrides={"task.domain_randomization": bad})
