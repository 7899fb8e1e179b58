"""Self-collision workload statistics of random-policy Humanoid states (numpy, from states the GPU stepped): per env the
number of link pairs whose bounding spheres overlap, the sphere tests those imply, and the sphere pairs in contact."""
import contextlib, io, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np
import torch
from omniisaacgymenvs_b200.utils.task_util import make_env

def qmat(q):
    w, x, y, z = q[..., 0], q[..., 1], q[..., 2], q[..., 3]
    return np.stack([np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)], -1),
                     np.stack([2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)], -1),
                     np.stack([2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], -1)], -2)

def fk(m, quat, q):
    N, L = q.shape[0], m.num_links
    R, p = [None] * L, [None] * L
    R[0] = qmat(quat); p[0] = np.zeros((N, 3))
    for i in range(1, L):
        pa, d, ax = int(m.parent[i]), int(m.dof_index[i]), m.joint_axis[i]
        Rj = R[pa] @ qmat(np.asarray(m.joint_quat[i]))[None]
        th = q[:, d]
        R[i] = Rj @ qmat(np.concatenate([np.cos(th / 2)[:, None], np.sin(th / 2)[:, None] * ax[None]], 1))
        p[i] = p[pa] + np.einsum('nij,j->ni', R[pa], m.joint_pos[i])
    return R, p

n = 2048
with contextlib.redirect_stdout(io.StringIO()):
    env, task = make_env("Humanoid", num_envs=n); env.reset()
dev = torch.device(task.device); v = task.get_robot(); m = v.model
for t in range(40):
    env.step(torch.rand(n, 21, device=dev) * 2 - 1)
    if t in (0, 9, 19, 39):
        quat = v.get_world_poses(clone=True)[1].cpu().numpy().astype(np.float64)
        q = v.get_joint_positions(clone=True).cpu().numpy().astype(np.float64)
        R, p = fk(m, quat, q)
        c = [p[int(l)] + np.einsum('nij,j->ni', R[int(l)], m.sphere_pos[k]) for k, l in enumerate(m.sphere_link)]
        nb = np.zeros(n); ns = np.zeros(n); nc = np.zeros(n)
        for la, lb in m.collision_pairs:
            ia = np.nonzero(m.sphere_link == la)[0]; ib = np.nonzero(m.sphere_link == lb)[0]
            ma = 0.5 * (c[ia[0]] + c[ia[-1]]); mb = 0.5 * (c[ib[0]] + c[ib[-1]])
            ra = max(np.linalg.norm(m.sphere_pos[i] - 0.5 * (m.sphere_pos[ia[0]] + m.sphere_pos[ia[-1]])) + m.sphere_radius[i] for i in ia)
            rb = max(np.linalg.norm(m.sphere_pos[i] - 0.5 * (m.sphere_pos[ib[0]] + m.sphere_pos[ib[-1]])) + m.sphere_radius[i] for i in ib)
            ov = np.linalg.norm(ma - mb, axis=1) < ra + rb
            nb += ov; ns += ov * len(ia) * len(ib)
            for i in ia:
                for j in ib:
                    nc += np.linalg.norm(c[i] - c[j], axis=1) < m.sphere_radius[i] + m.sphere_radius[j]
        print(f"step {t}: bounding overlaps/env mean {nb.mean():.1f} max {nb.max():.0f}; sphere tests mean {ns.mean():.0f} max {ns.max():.0f}; "
              f"sphere pairs in contact mean {nc.mean():.2f} max {nc.max():.0f}; envs with contact {np.mean(nc > 0):.2f}")
