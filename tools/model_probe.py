"""Model-level probes on the GPU (no oracle): (1) a robot made nearly rigid by stiff joint springs is set on its feet
and the foot sensors are summed against m*g; (2) random-policy episode lengths.  Prints what the tests then bound."""
import contextlib, io, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np
import torch
from omniisaacgymenvs_b200.utils.task_util import make_env

def standing(robot, height, stiff=3000.0):
    with contextlib.redirect_stdout(io.StringIO()):
        env, task = make_env(robot, num_envs=64)
    v = task.get_robot()
    m = v.model
    print(robot, "mass", m.total_mass(), "weight", m.total_mass() * 9.81)
    import ctypes as C
    from omniisaacgymenvs_b200 import _lib
    from omniisaacgymenvs_b200.robots.model import fill_desc
    m.stiffness[1:] = stiff; m.damping[1:] = 30.0
    desc = _lib.OigeModelDesc(); fill_desc(m, desc)
    lib = _lib.load()
    h = lib.oige_model_create(C.byref(desc))
    env.reset()
    dev = torch.device(task.device)
    n = 64
    pos, quat = v.get_world_poses(clone=True)
    pos[:, 2] = height
    q = torch.zeros(n, m.num_dof, device=dev) + task.initial_dof_pos.to(dev)
    # springs pull towards spring_ref = 0: use the q=0 pose
    q[:] = 0.0
    v.set_world_poses(pos, quat, torch.arange(n, device=dev)); v.set_velocities(torch.zeros(n, 6, device=dev), torch.arange(n, device=dev))
    v.set_joint_positions(q, torch.arange(n, device=dev)); v.set_joint_velocities(torch.zeros_like(q), torch.arange(n, device=dev))
    a = torch.zeros(n, m.num_dof, device=dev)
    st = v.state_struct
    for k in range(400):
        _lib.check(lib.oige_articulation_step_f32(h, C.byref(task._params), C.byref(st), a.data_ptr(), n, 2, 0.0083, _lib.current_stream_ptr(dev)), "k2")
        if k % 50 == 49:
            torch.cuda.synchronize()
            f = v._physics_view.get_force_sensor_forces()
            pz = v.get_world_poses(clone=True)[0][0, 2].item()
            print(k, "z", round(pz, 4), "sum |F|", float(f[0, :, :3].norm(dim=-1).sum()), "Fz-ish", f[0, :, :3].cpu().numpy().round(2).tolist(), "qmax", float(v.get_joint_positions()[0].abs().max()))

def episodes(robot, n=4096, steps=400):
    with contextlib.redirect_stdout(io.StringIO()):
        env, task = make_env(robot, num_envs=n); env.reset()
    dev = torch.device(task.device)
    g = torch.Generator(device=dev).manual_seed(1)
    lens = []
    for t in range(steps):
        _, _, done, _ = env.step(torch.rand(n, task.num_actions, device=dev, generator=g) * 2 - 1)
        d = done.nonzero().flatten()
        if len(d): lens.append(task.progress_buf[d].cpu().numpy())
    lens = np.concatenate(lens)
    print(robot, "episodes", len(lens), "len min/median/mean/p95/max", lens.min(), np.median(lens), lens.mean().round(2), np.percentile(lens, 95), lens.max())
    print("  histogram (10-step bins up to 100):", np.histogram(lens, bins=list(range(0, 101, 10)) + [10000])[0].tolist())

if __name__ == "__main__":
    standing("Humanoid", 1.36)
    standing("Ant", 0.6)
    episodes("Humanoid"); episodes("Ant")
