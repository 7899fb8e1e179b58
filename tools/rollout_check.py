"""Long random-policy rollout sanity (GPU box): finiteness, parked (non-finite-guard) envs, height / speed stats."""
import os, sys, io, contextlib
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from omniisaacgymenvs_b200.utils.task_util import make_env
for robot, n in (("Humanoid", 16384), ("Ant", 16384), ("Humanoid", 2048), ("Ant", 2048)):
    with contextlib.redirect_stdout(io.StringIO()):
        env, task = make_env(robot, num_envs=n); env.reset()
    v = task.get_robot(); parked = 0; maxv = 0.0; maxqd = 0.0
    for step in range(1000):
        a = torch.rand(n, task.num_actions, device=task.device) * 2 - 1
        obs, rew, done, _ = env.step(a)
        if step % 50 == 0:
            parked += int((v.root_pos[:, 2] == -1.0).sum())
            maxv = max(maxv, float(v.root_vel.abs().max())); maxqd = max(maxqd, float(v.dof_vel.abs().max()))
            assert torch.isfinite(obs["obs"]).all() and torch.isfinite(rew).all()
    st = task.read_episode_stats()
    print(f"{robot} n={n}: parked(sampled)={parked} max|root vel|={maxv:.1f} max|qd|={maxqd:.1f} mean z={float(v.root_pos[:,2].mean()):.3f} "
          f"episodes={int(st[1])} mean len={float(st[2]/max(st[1],1)):.1f} mean rew/step={float(st[0]/st[3]):.3f}")
