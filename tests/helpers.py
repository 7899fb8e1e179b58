"""Shared test helpers: drive the CUDA path through the C ABI and compare it with the oracle."""
import ctypes as C
import math
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from omniisaacgymenvs_b200 import _lib  # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")
RTOL = 1e-5  # north_star: obs/reward within 1e-5 relative error in fp32
ATOL = 1e-5  # absolute floor for O(1) quantities (angles, projections, rewards near 0)


def params_from_spec(spec, dof_lo, dof_hi, progress_increment=1, clip_obs=None, walk_uses_z=1, closed_form_angles=1):
    p = _lib.OigeTaskParams()
    p.num_dof, p.num_sensors, p.num_obs = spec.num_dof, spec.num_sensors, spec.num_obs
    p.limit_cost_kind = _lib.LIMIT_COST_ANT if spec.name == "Ant" else _lib.LIMIT_COST_HUMANOID
    p.walk_target_uses_z = walk_uses_z
    p.progress_increment = progress_increment
    p.closed_form_angles = closed_form_angles
    p.dt = spec.dt
    p.dof_vel_scale = spec.dof_vel_scale
    p.angular_velocity_scale = spec.angular_velocity_scale
    p.contact_force_scale = spec.contact_force_scale
    p.heading_weight = spec.heading_weight
    p.up_weight = spec.up_weight
    p.actions_cost_scale = spec.actions_cost_scale
    p.energy_cost_scale = spec.energy_cost_scale
    p.joints_at_limit_cost_scale = spec.joints_at_limit_cost_scale
    p.death_cost = spec.death_cost
    p.termination_height = spec.termination_height
    p.alive_reward_scale = spec.alive_reward_scale
    p.max_episode_length = float(spec.max_episode_length)
    p.clip_obs = float(spec.clip_obs if clip_obs is None else clip_obs)
    p.power_scale = spec.power_scale
    for k, v in enumerate([1000.0, 0.0, 0.0]):
        p.targets[k] = v
    for k, v in enumerate([1.0, 0.0, 0.0, 0.0]):
        p.inv_start_rot[k] = v
    for k, v in enumerate([1.0, 0.0, 0.0]):
        p.basis_vec0[k] = v
    for k, v in enumerate([0.0, 0.0, 1.0]):
        p.basis_vec1[k] = v
    mer = spec.motor_effort_ratio().tolist()
    for k in range(spec.num_dof):
        p.dof_limits_lower[k] = float(dof_lo[k])
        p.dof_limits_upper[k] = float(dof_hi[k])
        p.motor_effort_ratio[k] = mer[k]
        p.joint_gears[k] = spec.joint_gears[k]
    return p


class DeviceCase:
    """Uploads one oracle-style state dict and owns the K1 output buffers."""

    def __init__(self, spec, s, device="cuda:0", with_stats=False):
        self.spec = spec
        self.n = s["pos"].shape[0]
        dev = torch.device(device)
        up = lambda t: t.detach().clone().contiguous().to(dev)
        self.pos, self.quat, self.vel = up(s["pos"]), up(s["quat"]), up(s["vel"])
        self.dof_pos, self.dof_vel = up(s["dof_pos"]), up(s["dof_vel"])
        self.sensors, self.actions = up(s["sensors"]), up(s["actions"])
        self.potentials = up(s["potentials"])
        self.reset_buf, self.progress_buf = up(s["reset_buf"]), up(s["progress_buf"])
        n = self.n
        f = dict(device=dev, dtype=torch.float32)
        self.obs = torch.full((n, spec.num_obs), float("nan"), **f)
        self.rew = torch.full((n,), float("nan"), **f)
        self.prev_potentials = torch.full((n,), float("nan"), **f)
        self.up_vec = torch.full((n, 3), float("nan"), **f)
        self.heading_vec = torch.full((n, 3), float("nan"), **f)
        self.stats = torch.zeros((_lib.STATS_SLOTS, 4), device=dev, dtype=torch.float64) if with_stats else None
        self.dev = dev

    def structs(self):
        st = _lib.OigeState(self.pos.data_ptr(), self.quat.data_ptr(), self.vel.data_ptr(), self.dof_pos.data_ptr(),
                            self.dof_vel.data_ptr(), self.sensors.data_ptr(), None)
        bf = _lib.OigeTaskBuffers(self.obs.data_ptr(), self.rew.data_ptr(), self.reset_buf.data_ptr(),
                                  self.progress_buf.data_ptr(), self.potentials.data_ptr(),
                                  self.prev_potentials.data_ptr(), self.up_vec.data_ptr(),
                                  self.heading_vec.data_ptr(), self.actions.data_ptr(),
                                  self.stats.data_ptr() if self.stats is not None else None, None)
        return st, bf

    def outputs(self):
        torch.cuda.synchronize(self.dev)
        return {"obs": self.obs.cpu(), "rew": self.rew.cpu(), "reset_buf": self.reset_buf.cpu(),
                "progress_buf": self.progress_buf.cpu(), "potentials": self.potentials.cpu(),
                "prev_potentials": self.prev_potentials.cpu(), "up_vec": self.up_vec.cpu(),
                "heading_vec": self.heading_vec.cpu()}


def run_k1(spec, s, device="cuda:0", tile=0, clip_obs=None, progress_increment=1, with_stats=False,
           closed_form_angles=1):
    """One fused obs/reward/reset pass through the C ABI on a copy of state dict `s`."""
    lib = _lib.load()
    case = DeviceCase(spec, s, device, with_stats=with_stats)
    p = params_from_spec(spec, s["dof_lo"], s["dof_hi"], progress_increment=progress_increment, clip_obs=clip_obs,
                         closed_form_angles=closed_form_angles)
    st, bf = case.structs()
    with torch.cuda.device(case.dev):
        stream = _lib.current_stream_ptr(case.dev)
        if tile:
            rc = lib.oige_obs_reward_reset_tile_f32(C.byref(p), C.byref(st), C.byref(bf), case.n, tile, stream)
        else:
            rc = lib.oige_obs_reward_reset_f32(C.byref(p), C.byref(st), C.byref(bf), case.n, stream)
    _lib.check(rc, "oige_obs_reward_reset_f32")
    out = case.outputs()
    if with_stats:
        out["stats"] = case.stats.sum(0).cpu()
    return out


def _circ_diff(a, b):
    d = (a - b).abs()
    return torch.minimum(d, (2 * math.pi - d).abs())


REW_ATOL = 1e-6  # measured (tests/k1_deviation.py, profiles/r2_k1_deviation.txt): max |d rew| 1.5e-5 at |rew| ~ 1e2,
REW_RTOL = 1e-6  # i.e. 2.4e-7 relative -- the per-dof sums are ordered differently from torch's vectorised reduction
ANGLE_TOL = 1.5e-6  # measured 6.0e-7 (closed-form wrap) / 7.2e-7 (literal chain) on the circle: 1-3 ulp at |x| ~ pi


def assert_k1_matches(got, ref, what=""):
    """The parity bar (north_star: 1e-5 relative in fp32, reset indices bit-exact), with CONSTANT tolerances taken
    from the measured deviation table: reset/progress/potentials/prev_potentials/height bit-exact; reward within
    1e-6 + 1e-6*|rew| (10x inside the bar); angle columns within 1.5e-6 on the circle; other floats within 1e-5
    relative (measured: bit-identical at 4096 envs; the bound covers torch's scalar-tail kernels at ragged N)."""
    assert torch.equal(got["reset_buf"], ref["reset_buf"]), f"{what}: reset_buf not bit-exact"
    assert torch.equal(got["progress_buf"], ref["progress_buf"]), f"{what}: progress_buf not bit-exact"
    assert torch.equal(got["prev_potentials"], ref["prev_potentials"]), f"{what}: prev_potentials must be a copy"
    assert torch.equal(got["potentials"], ref["potentials"]), (
        f"{what}: potentials not bit-exact (max |d| {float((got['potentials'] - ref['potentials']).abs().max())})")
    go, ro = got["obs"], ref["obs"]
    assert torch.equal(go[:, 0], ro[:, 0]), f"{what}: obs[:,0] (height) must be a bit-exact copy"
    ang_cols = [7, 8, 9]
    other = [c for c in range(go.shape[1]) if c not in ang_cols]
    torch.testing.assert_close(go[:, other], ro[:, other], rtol=RTOL, atol=ATOL, msg=f"{what}: obs")
    ad = _circ_diff(go[:, ang_cols], ro[:, ang_cols])
    assert float(ad.max()) <= ANGLE_TOL, f"{what}: angle columns differ by {float(ad.max())}"
    torch.testing.assert_close(got["up_vec"], ref["up_vec"], rtol=RTOL, atol=ATOL, msg=f"{what}: up_vec")
    torch.testing.assert_close(got["heading_vec"], ref["heading_vec"], rtol=RTOL, atol=ATOL, msg=f"{what}: heading_vec")
    tol = REW_ATOL + REW_RTOL * ref["rew"].abs()
    bad = (got["rew"] - ref["rew"]).abs() > tol
    # a reward branch (up>0.93 / heading>0.8 / ant q>0.99 count) may legitimately flip when its input sits within
    # one ulp of the threshold; such rows must be vanishingly rare
    assert int(bad.sum()) <= max(0, got["rew"].numel() // 100000), (
        f"{what}: reward mismatch on {int(bad.sum())} rows, max err {float((got['rew'] - ref['rew']).abs().max())}")


def load_golden(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    s = {k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("in_")}
    out = {k[4:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("out_")}
    return s, out


def golden_names():
    return sorted(f[:-4] for f in os.listdir(GOLDEN_DIR) if f.endswith(".npz") and not f.startswith("episode_"))
