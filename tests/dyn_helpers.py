"""Helpers for the K2 (articulation step) parity tests: run the CUDA kernel through the C ABI and
the float64 CPU oracle on the same state."""
import ctypes as C

import numpy as np
import torch

from omniisaacgymenvs_b200 import _lib
from omniisaacgymenvs_b200.robots.model import build_model, fill_desc
from oracle import dynamics as dyn
from oracle import locomotion_oracle as lo


def random_state(model, n, seed, height=None, qd_scale=1.0, root_vel_scale=0.5, tilt=0.3):
    """Random but physically sane state: joints inside their limits, moderate velocities."""
    rng = np.random.default_rng(seed)
    lim = model.dof_limits().astype(np.float64)
    D = model.num_dof
    q = lim[:, 0] + (0.15 + 0.7 * rng.random((n, D))) * (lim[:, 1] - lim[:, 0])
    qd = qd_scale * rng.normal(size=(n, D))
    pos = np.zeros((n, 3))
    pos[:, :2] = rng.normal(size=(n, 2)) * 3.0
    pos[:, 2] = height if height is not None else model.spawn_height
    axis = rng.normal(size=(n, 3))
    axis /= np.linalg.norm(axis, axis=1, keepdims=True)
    ang = tilt * rng.normal(size=(n, 1))
    quat = np.concatenate([np.cos(ang / 2), np.sin(ang / 2) * axis], axis=1)
    vel = root_vel_scale * rng.normal(size=(n, 6))
    actions = rng.uniform(-1, 1, size=(n, D))
    f32 = lambda a: np.ascontiguousarray(a.astype(np.float32))
    return {"pos": f32(pos), "quat": f32(quat), "vel": f32(vel), "q": f32(q), "qd": f32(qd), "actions": f32(actions)}


def unit_params(model, spec):
    p = _lib.OigeTaskParams()
    p.num_dof, p.num_sensors = model.num_dof, model.num_sensors
    p.num_obs = 12 + 3 * model.num_dof + 6 * model.num_sensors
    p.power_scale = spec.power_scale
    for k in range(model.num_dof):
        p.joint_gears[k] = float(model.joint_gears[k])
    return p


def run_gpu(cuda_lib, model, spec, st, substeps, dt, device="cuda:0", calls=1):
    dev = torch.device(device)
    desc = _lib.OigeModelDesc()
    fill_desc(model, desc)
    with torch.cuda.device(dev):
        handle = cuda_lib.oige_model_create(C.byref(desc))
    assert handle, cuda_lib.oige_last_error()
    t = {k: torch.from_numpy(v.copy()).to(dev) for k, v in st.items()}
    n = st["pos"].shape[0]
    sens = torch.zeros((n, model.num_sensors, 6), device=dev)
    state = _lib.OigeState(t["pos"].data_ptr(), t["quat"].data_ptr(), t["vel"].data_ptr(), t["q"].data_ptr(),
                           t["qd"].data_ptr(), sens.data_ptr(), t["targets"].data_ptr() if "targets" in t else None)
    p = unit_params(model, spec)
    for _ in range(calls):
        rc = cuda_lib.oige_articulation_step_f32(handle, C.byref(p), C.byref(state), t["actions"].data_ptr(), n,
                                                 substeps, dt, _lib.current_stream_ptr(dev))
        _lib.check(rc, "oige_articulation_step_f32")
    torch.cuda.synchronize(dev)
    cuda_lib.oige_model_destroy(handle)
    out = {k: v.cpu().numpy().astype(np.float64) for k, v in t.items() if k not in ("actions", "targets")}
    out["sensors"] = sens.cpu().numpy().astype(np.float64)
    return out


def run_oracle(model, spec, st, substeps, dt, calls=1, threads=8):
    desc = dyn.make_desc(model)
    d = {k: np.ascontiguousarray(v.astype(np.float64)) for k, v in st.items()}
    tg = d.get("targets")
    gears = model.joint_gears.astype(np.float32)
    # efforts formed in fp32 exactly as the kernel / reference do: (a * gear) * power_scale
    eff = ((st["actions"] * gears).astype(np.float32) * np.float32(spec.power_scale)).astype(np.float64)
    eff = np.ascontiguousarray(eff)
    n = st["pos"].shape[0]
    sens = np.zeros((n, model.num_sensors, 6))
    for _ in range(calls):
        bad = dyn.step(desc, d["pos"], d["quat"], d["vel"], d["q"], d["qd"], eff, sens, substeps, dt, threads, targets=tg)
        assert bad == 0, f"oracle: {bad} envs with a non-SPD system"
    out = {k: d[k] for k in ("pos", "quat", "vel", "q", "qd")}
    out["sensors"] = sens
    return out


def specs(robot, **overrides):
    """Model + task constants.  Self-collision (Humanoid) is integrated by the role-split kernel only, so the
    warp-per-env / thread-per-env kernels (OIGE_K2_MODE=warp|thread) are tested on the model without it."""
    import os

    if os.environ.get("OIGE_K2_MODE", "") in ("warp", "thread"):
        overrides.setdefault("enable_self_collisions", False)
    return build_model(robot, **overrides), lo.SPECS[robot]


def self_contact_states(m, n, seed, min_depth=0.01, height=50.0, qd_scale=0.3):
    """Random joint configurations (fp32-exact) in which at least one self-collision sphere pair is in contact."""
    rng = np.random.default_rng(seed)
    lim = m.dof_limits().astype(np.float64)
    out = []
    while len(out) < n:
        q = (lim[:, 0] + rng.random(m.num_dof) * (lim[:, 1] - lim[:, 0])).astype(np.float32).astype(np.float64)
        s = dict(pos=np.array([[0.0, 0.0, height]]), quat=np.array([[1.0, 0, 0, 0]]), q=q[None])
        if self_contact_depths(m, s).max() > min_depth:
            out.append(q)
    q = np.array(out)
    qd = (qd_scale * rng.normal(size=q.shape)).astype(np.float32).astype(np.float64)
    return dict(pos=np.tile([[0.0, 0.0, height]], (n, 1)), quat=np.tile([[1.0, 0, 0, 0]], (n, 1)), vel=np.zeros((n, 6)),
                q=q, qd=qd)


def link_frames(model, s, e=0):
    """(R[L], p[L]) of env e: float64 numpy forward kinematics, hinge + prismatic."""
    def qmat(q):
        w, x, y, z = q
        return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                         [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                         [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])

    L = model.num_links
    R, p = [None] * L, [None] * L
    R[0], p[0] = qmat(np.asarray(s["quat"][e], float)), np.array(s["pos"][e], float)
    for i in range(1, L):
        pa, d = int(model.parent[i]), int(model.dof_index[i])
        ax = model.joint_axis[i]
        Rj = R[pa] @ qmat(model.joint_quat[i])
        if model.joint_type[i] == 1:
            R[i] = Rj
            p[i] = p[pa] + R[pa] @ model.joint_pos[i] + (Rj @ ax) * float(s["q"][e, d])
        else:
            th = float(s["q"][e, d])
            R[i] = Rj @ qmat(np.concatenate([[np.cos(th / 2)], np.sin(th / 2) * ax]))
            p[i] = p[pa] + R[pa] @ model.joint_pos[i]
    return R, p


def _segment_distance(P1, Q1, P2, Q2):
    d1, d2, r = Q1 - P1, Q2 - P2, P1 - P2
    a, e, f, c, b = d1 @ d1, d2 @ d2, d2 @ r, d1 @ r, d1 @ d2
    if a <= 1e-12 and e <= 1e-12:
        s = t = 0.0
    elif a <= 1e-12:
        s, t = 0.0, min(max(f / e, 0.0), 1.0)
    elif e <= 1e-12:
        t, s = 0.0, min(max(-c / a, 0.0), 1.0)
    else:
        den = a * e - b * b
        s = min(max((b * f - c * e) / den, 0.0), 1.0) if den > 1e-9 * a * e else 0.0
        t = (b * s + f) / e
        if t < 0:
            t, s = 0.0, min(max(-c / a, 0.0), 1.0)
        elif t > 1:
            t, s = 1.0, min(max((b - c) / a, 0.0), 1.0)
    return float(np.linalg.norm((P1 + s * d1) - (P2 + t * d2)))


def self_contact_depths(model, s, e=0):
    """Penetration depth (> 0: in contact) of the capsules of every self-collision link pair of the model, env e."""
    R, p = link_frames(model, s, e)
    out = []
    for la, lb in model.collision_pairs:
        la, lb = int(la), int(lb)
        P1, Q1 = p[la] + R[la] @ model.capsule_a[la], p[la] + R[la] @ model.capsule_b[la]
        P2, Q2 = p[lb] + R[lb] @ model.capsule_a[lb], p[lb] + R[lb] @ model.capsule_b[lb]
        out.append(model.capsule_radius[la] + model.capsule_radius[lb] - _segment_distance(P1, Q1, P2, Q2))
    return np.array(out)


def link_coms(model, s):
    """Mass-weighted centre of mass of env 0 (float64 numpy forward kinematics, hinge + prismatic)."""
    def qmat(q):
        w, x, y, z = q
        return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                         [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                         [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])

    L = model.num_links
    R, p = [None] * L, [None] * L
    R[0], p[0] = qmat(s["quat"][0]), np.array(s["pos"][0], float)
    for i in range(1, L):
        pa, d = int(model.parent[i]), int(model.dof_index[i])
        ax = model.joint_axis[i]
        Rj = R[pa] @ qmat(model.joint_quat[i])
        if model.joint_type[i] == 1:
            R[i] = Rj
            p[i] = p[pa] + R[pa] @ model.joint_pos[i] + (Rj @ ax) * s["q"][0, d]
        else:
            th = s["q"][0, d]
            R[i] = Rj @ qmat(np.concatenate([[np.cos(th / 2)], np.sin(th / 2) * ax]))
            p[i] = p[pa] + R[pa] @ model.joint_pos[i]
    tot = sum(model.mass)
    return sum(model.mass[i] * (p[i] + R[i] @ model.com[i]) for i in range(L)) / tot
