"""CPU: the restated third-party helpers (oracle/isaac_math.py) against an independent float64
numpy / scipy restatement -- the only pin available for them (source absent)."""
import math

import numpy as np
import torch
from scipy.spatial.transform import Rotation

from oracle import isaac_math as im


def _rand_quat(n, seed):
    g = torch.Generator().manual_seed(seed)
    q = torch.randn(n, 4, generator=g)
    return q / q.norm(dim=-1, keepdim=True)


def _scipy_rot(q_wxyz):
    q = q_wxyz.double().numpy()
    return Rotation.from_quat(np.stack([q[:, 1], q[:, 2], q[:, 3], q[:, 0]], -1))


def test_quat_mul_matches_hamilton_product():
    a, b = _rand_quat(512, 0), _rand_quat(512, 1)
    got = im.quat_mul(a, b).double().numpy()
    ref = (_scipy_rot(a) * _scipy_rot(b)).as_quat()
    ref = np.stack([ref[:, 3], ref[:, 0], ref[:, 1], ref[:, 2]], -1)
    sign = np.sign((got * ref).sum(-1, keepdims=True))
    assert np.abs(got - sign * ref).max() < 2e-6


def test_quat_conjugate_and_identity():
    a = _rand_quat(64, 2)
    e = im.quat_mul(a, im.quat_conjugate(a))
    assert torch.allclose(e, torch.tensor([1.0, 0, 0, 0]).expand_as(e), atol=1e-6)


def test_quat_rotate_and_inverse():
    q = _rand_quat(512, 3)
    v = torch.randn(512, 3, generator=torch.Generator().manual_seed(4))
    ref = _scipy_rot(q).apply(v.double().numpy())
    assert np.abs(im.quat_rotate(q, v).double().numpy() - ref).max() < 5e-6
    back = im.quat_rotate_inverse(q, im.quat_rotate(q, v))
    assert torch.allclose(back, v, atol=5e-6)


def test_get_euler_xyz_matches_scipy_and_wraps_to_0_2pi():
    q = _rand_quat(2048, 5)
    roll, pitch, yaw = im.get_euler_xyz(q)
    for ang in (roll, pitch, yaw):
        assert float(ang.min()) >= 0.0 and float(ang.max()) < 2 * math.pi + 1e-6
    e = _scipy_rot(q).as_euler("xyz")  # extrinsic xyz == roll, pitch, yaw
    wrap = lambda x: np.mod(x, 2 * math.pi)
    circ = lambda a, b: np.minimum(np.abs(a - b), 2 * math.pi - np.abs(a - b))
    keep = np.abs(np.abs(e[:, 1]) - math.pi / 2) > 1e-2  # away from gimbal lock
    assert circ(roll.double().numpy(), wrap(e[:, 0]))[keep].max() < 1e-4
    assert circ(pitch.double().numpy(), wrap(e[:, 1]))[keep].max() < 1e-4
    assert circ(yaw.double().numpy(), wrap(e[:, 2]))[keep].max() < 1e-4


def test_gimbal_branch_returns_half_pi():
    r2 = math.sqrt(0.5)
    q = torch.tensor([[r2, 0.0, r2, 0.0], [r2, 0.0, -r2, 0.0]])
    _, pitch, _ = im.get_euler_xyz(q)
    assert abs(float(pitch[0]) - math.pi / 2) < 1e-3
    assert abs(float(pitch[1]) - (2 * math.pi - math.pi / 2)) < 1e-3


def test_compute_heading_and_up_identity_pose():
    n = 4
    q = torch.tensor([1.0, 0, 0, 0]).repeat(n, 1)
    to_target = torch.tensor([3.0, 4.0, 0.0]).repeat(n, 1)
    v0 = torch.tensor([1.0, 0, 0]).repeat(n, 1)
    v1 = torch.tensor([0.0, 0, 1]).repeat(n, 1)
    tq, up_proj, heading_proj, up_vec, heading_vec = im.compute_heading_and_up(q, q.clone(), to_target, v0, v1, 2)
    assert torch.allclose(up_proj, torch.ones(n))
    assert torch.allclose(heading_proj, torch.full((n,), 0.6))
    assert torch.allclose(tq, q)


def test_compute_rot_uses_z_quirk():
    q = torch.tensor([[1.0, 0, 0, 0]])
    pos = torch.tensor([[0.0, 5.0, 2.0]])
    tgt = torch.tensor([[1000.0, 0.0, 0.0]])
    z = torch.zeros(1, 3)
    *_, angle = im.compute_rot(q, z, z, tgt, pos)
    assert abs(float(angle) - math.atan2(-2.0, 1000.0)) < 1e-7  # dz, not dy


def test_unscale_clamp_rand():
    lo, hi = torch.tensor([-1.0, 0.0]), torch.tensor([1.0, 2.0])
    assert torch.allclose(im.unscale(torch.tensor([[1.0, 0.0]]), lo, hi), torch.tensor([[1.0, -1.0]]))
    assert torch.equal(im.tensor_clamp(torch.tensor([[-5.0, 5.0]]), lo, hi), torch.tensor([[-1.0, 2.0]]))
    torch.manual_seed(0)
    r = im.torch_rand_float(-0.2, 0.2, (1000, 3), "cpu")
    assert float(r.min()) >= -0.2 and float(r.max()) <= 0.2


def test_half_angle_polynomials_absolute_error_up_to_the_cut_over():
    """csrc/articulation_step.cu `sincos_half_angle`: the two Taylor chains K2 evaluates instead of sincosf for
    |x| <= 1.45 rad, restated here with the same coefficients and the same fused multiply-adds (float64 product + sum,
    rounded once to float32).  The bound that matters for a unit quaternion is ABSOLUTE: <= 1.1e-7 (< 2 ulp of 1.0)
    inside the cut-over; between 1.45 and the old cut-over (1.75 rad) the error passes 1.3e-7, which is why the
    cut-over was lowered -- there the kernel calls sincosf."""
    f32 = np.float32

    def fma(a, b, c):
        return (a.astype(np.float64) * b.astype(np.float64) + np.asarray(c, np.float64)).astype(f32)

    def chains(x):
        x2 = (x * x).astype(f32)
        ps = fma(x2, f32(1.6059044e-10), f32(-2.5052108e-8))
        for c in (2.7557319e-6, -1.9841270e-4, 8.3333333e-3, -1.6666667e-1):
            ps = fma(ps, x2, f32(c))
        s = fma((ps * x2).astype(f32), x, x)
        pc = fma(x2, f32(2.0876757e-9), f32(-2.7557319e-7))
        for c in (2.4801587e-5, -1.3888889e-3, 4.1666667e-2, -0.5):
            pc = fma(pc, x2, f32(c))
        return s, fma(pc, x2, f32(1.0))

    x = np.linspace(0.0, 1.45, 200001).astype(f32)
    x = np.concatenate([x, -x])
    s, c = chains(x)
    xd = x.astype(np.float64)
    assert np.abs(s - np.sin(xd)).max() <= 1.1e-7 and np.abs(c - np.cos(xd)).max() <= 1.1e-7
    assert np.abs(s.astype(np.float64) ** 2 + c.astype(np.float64) ** 2 - 1.0).max() <= 2.5e-7
    far = np.linspace(1.45, 1.75, 20001).astype(f32)
    fs, fc = chains(far)
    fd = far.astype(np.float64)
    assert max(np.abs(fs - np.sin(fd)).max(), np.abs(fc - np.cos(fd)).max()) > 1.3e-7
