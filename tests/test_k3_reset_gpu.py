"""GPU: K3 (device-side reset) through the C ABI against the oracle's reset_idx restatement
(tasks/shared/locomotion.py:116-145).  With supplied uniforms (the reference's torch.rand draws)
every written value must be bit-exact; with in-kernel Philox the draws must equal the numpy
Philox4x32-10 restatement (oracle/philox.py) bit for bit and be sharding-invariant."""
import ctypes as C

import numpy as np
import pytest
import torch

import helpers
from omniisaacgymenvs_b200 import _lib
from oracle import locomotion_oracle as lo
from oracle import philox

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _setup(spec, n, seed):
    g = torch.Generator().manual_seed(seed)
    s = lo.synthetic_state(spec, n, seed)
    s["prev_potentials"] = s["potentials"].clone()
    s["initial_root_pos"] = torch.randn(n, 3, generator=g) * 20
    s["initial_root_pos"][:, 2] = 1.34
    s["initial_root_rot"] = torch.tensor([1.0, 0, 0, 0]).repeat(n, 1)
    s["initial_dof_pos"] = torch.zeros(n, spec.num_dof)
    s["reset_buf"] = (torch.rand(n, generator=g) < 0.3).long()
    return s


def _device_run(spec, s, fn):
    dev = torch.device(DEV)
    case = helpers.DeviceCase(spec, s, DEV)
    case.prev_potentials = s["prev_potentials"].clone().to(dev)
    init = [s[k].clone().contiguous().to(dev) for k in ("initial_root_pos", "initial_root_rot", "initial_dof_pos")]
    st, bf = case.structs()
    ri = _lib.OigeResetInit(*[t.data_ptr() for t in init])
    p = helpers.params_from_spec(spec, s["dof_lo"], s["dof_hi"])
    fn(p, st, bf, ri, dev)
    torch.cuda.synchronize(dev)
    return {"pos": case.pos.cpu(), "quat": case.quat.cpu(), "vel": case.vel.cpu(), "dof_pos": case.dof_pos.cpu(),
            "dof_vel": case.dof_vel.cpu(), "potentials": case.potentials.cpu(),
            "prev_potentials": case.prev_potentials.cpu(), "reset_buf": case.reset_buf.cpu(),
            "progress_buf": case.progress_buf.cpu()}


@pytest.mark.parametrize("robot", ["Humanoid", "Ant"])
@pytest.mark.parametrize("n", [1, 33, 1000, 4096])
def test_masked_reset_with_supplied_uniforms_is_bit_exact(cuda_lib, robot, n):
    spec = lo.SPECS[robot]
    s = _setup(spec, n, n)
    g = torch.Generator().manual_seed(99)
    rp, rv = torch.rand(n, spec.num_dof, generator=g), torch.rand(n, spec.num_dof, generator=g)
    ids = s["reset_buf"].nonzero(as_tuple=False).squeeze(-1)
    ref = {k: v.clone() for k, v in s.items()}
    lo.reset_idx(spec, ref, ids, rp[ids], rv[ids])

    def call(p, st, bf, ri, dev):
        drp, drv = rp.to(dev), rv.to(dev)
        _lib.check(cuda_lib.oige_reset_masked_f32(C.byref(p), C.byref(st), C.byref(bf), C.byref(ri), drp.data_ptr(),
                                                  drv.data_ptr(), 0, 0, 0, n, _lib.current_stream_ptr(dev)), "k3")
        torch.cuda.synchronize(dev)

    got = _device_run(spec, s, call)
    for k in got:
        assert torch.equal(got[k], ref[k]), f"{k} not bit-exact"
    assert int(got["reset_buf"].sum()) == 0


@pytest.mark.parametrize("robot", ["Humanoid", "Ant"])
def test_indexed_reset_matches_reference_calling_form(cuda_lib, robot):
    spec = lo.SPECS[robot]
    n = 777
    s = _setup(spec, n, 3)
    ids = torch.tensor([5, 0, 776, 123, 400], dtype=torch.int64)
    g = torch.Generator().manual_seed(5)
    rp, rv = torch.rand(len(ids), spec.num_dof, generator=g), torch.rand(len(ids), spec.num_dof, generator=g)
    ref = {k: v.clone() for k, v in s.items()}
    lo.reset_idx(spec, ref, ids, rp, rv)

    def call(p, st, bf, ri, dev):
        dids, drp, drv = ids.to(dev), rp.to(dev), rv.to(dev)
        _lib.check(cuda_lib.oige_reset_idx_f32(C.byref(p), C.byref(st), C.byref(bf), C.byref(ri), dids.data_ptr(),
                                               len(ids), drp.data_ptr(), drv.data_ptr(), 0, 0, 0,
                                               _lib.current_stream_ptr(dev)), "k3 idx")
        torch.cuda.synchronize(dev)

    got = _device_run(spec, s, call)
    for k in got:
        assert torch.equal(got[k], ref[k]), f"{k} not bit-exact"


def test_philox_reset_is_bit_exact_and_sharding_invariant(cuda_lib):
    spec = lo.HUMANOID
    n, seed, counter = 512, 0x1234ABCD5678, 77
    s = _setup(spec, n, 8)
    s["reset_buf"][:] = 1

    def make_call(offset, count, lo_row):
        def call(p, st, bf, ri, dev):
            _lib.check(cuda_lib.oige_reset_masked_f32(C.byref(p), C.byref(st), C.byref(bf), C.byref(ri), None, None,
                                                      seed, counter, offset, count, _lib.current_stream_ptr(dev)), "k3")
        return call

    full = _device_run(spec, s, make_call(0, n, 0))
    rp = philox.reset_uniforms(seed, counter, np.arange(n), spec.num_dof, 0)
    rv = philox.reset_uniforms(seed, counter, np.arange(n), spec.num_dof, 1)
    ref = {k: v.clone() for k, v in s.items()}
    lo.reset_idx(spec, ref, torch.arange(n), torch.from_numpy(rp), torch.from_numpy(rv))
    for k in full:
        assert torch.equal(full[k], ref[k]), f"{k}: in-kernel Philox differs from the numpy restatement"
    # shard 1 of 2 run alone (rows 256:512, env_id_offset 256) == rows of the full run
    half = {k: (v[256:].clone() if v.shape[:1] == (n,) else v) for k, v in s.items()}
    part = _device_run(spec, half, make_call(256, 256, 256))
    for k in part:
        assert torch.equal(part[k], full[k][256:]), f"{k}: result depends on sharding"


def test_reset_noise_statistics(cuda_lib):
    spec = lo.ANT
    n = 65536
    s = _setup(spec, n, 1)
    s["reset_buf"][:] = 1
    s["dof_lo"] = torch.full((8,), -10.0)
    s["dof_hi"] = torch.full((8,), 10.0)

    def call(p, st, bf, ri, dev):
        _lib.check(cuda_lib.oige_reset_masked_f32(C.byref(p), C.byref(st), C.byref(bf), C.byref(ri), None, None, 1, 1,
                                                  0, n, _lib.current_stream_ptr(dev)), "k3")

    got = _device_run(spec, s, call)
    q, qd = got["dof_pos"], got["dof_vel"]
    assert float(q.min()) >= -0.2 and float(q.max()) <= 0.2 and abs(float(q.mean())) < 2e-3
    assert float(qd.min()) >= -0.1 and float(qd.max()) <= 0.1 and abs(float(qd.mean())) < 1e-3
    assert abs(float(q.std()) - 0.4 / 12 ** 0.5) < 2e-3
    assert abs(float(torch.corrcoef(torch.stack([q[:, 0], qd[:, 0]]))[0, 1])) < 0.02
