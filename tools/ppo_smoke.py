"""Behaviour-level check of the stand-in physics: a minimal PPO (plain torch, no rl_games -- it is not
installable here) on the stock Ant / Humanoid task through `VecEnvRLGames.step`, CUDA-graph mode.  If the mechanical model
were implausible (no traction, energy leaks, exploding contacts) the return would not rise.  Hyper-parameters follow the
reference's cfg/train/AntPPO.yaml where they apply (gamma 0.99, tau 0.95, e_clip 0.2, horizon 16, 4 mini-epochs, lr 3e-4,
normalised observations / advantages / values, MLP 256-128-64 ELU).

    python tools/ppo_smoke.py [Ant|Humanoid] [iterations] [num_envs]

Prints the mean reward per step and the mean episode length every 10 iterations; `run()` returns the curve."""
import contextlib
import io
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.nn as nn

from omniisaacgymenvs_b200.utils.task_util import make_env


class RunningNorm:
    def __init__(self, shape, dev):
        self.mean = torch.zeros(shape, device=dev); self.var = torch.ones(shape, device=dev); self.count = 1e-4

    def update(self, x):
        b_mean, b_var, n = x.mean(0), x.var(0, unbiased=False), x.shape[0]
        d = b_mean - self.mean; tot = self.count + n
        self.mean = self.mean + d * n / tot
        self.var = (self.var * self.count + b_var * n + d * d * self.count * n / tot) / tot
        self.count = tot

    def __call__(self, x):
        return ((x - self.mean) / torch.sqrt(self.var + 1e-5)).clamp(-5, 5)


class Policy(nn.Module):
    def __init__(self, n_obs, n_act):
        super().__init__()
        self.body = nn.Sequential(nn.Linear(n_obs, 256), nn.ELU(), nn.Linear(256, 128), nn.ELU(), nn.Linear(128, 64), nn.ELU())
        self.mu = nn.Linear(64, n_act); self.v = nn.Linear(64, 1)
        self.log_std = nn.Parameter(torch.zeros(n_act))

    def forward(self, obs):
        h = self.body(obs)
        return self.mu(h), self.log_std.expand(obs.shape[0], -1), self.v(h).squeeze(-1)


def run(robot="Ant", iterations=150, num_envs=4096, horizon=16, seed=0, verbose=True):
    torch.manual_seed(seed)
    with contextlib.redirect_stdout(io.StringIO()):
        env, task = make_env(robot, num_envs=num_envs, seed=seed)
        obs = env.reset()["obs"].clone()
        env.enable_cuda_graph()
    dev = torch.device(task.device)
    n_obs, n_act = task.num_observations, task.num_actions
    pi = Policy(n_obs, n_act).to(dev)
    opt = torch.optim.Adam(pi.parameters(), lr=3e-4)
    obs_norm, ret_norm = RunningNorm((n_obs,), dev), RunningNorm((), dev)
    gamma, lam, clip, epochs, minibatches = 0.99, 0.95, 0.2, 4, 4
    curve, t0 = [], time.time()
    for it in range(iterations):
        O = torch.empty(horizon, num_envs, n_obs, device=dev); A = torch.empty(horizon, num_envs, n_act, device=dev)
        LP = torch.empty(horizon, num_envs, device=dev); V = torch.empty(horizon + 1, num_envs, device=dev)
        R = torch.empty(horizon, num_envs, device=dev); D = torch.empty(horizon, num_envs, device=dev)
        task.read_episode_stats()
        with torch.no_grad():
            for t in range(horizon):
                obs_norm.update(obs)
                mu, log_std, v = pi(obs_norm(obs))
                a = mu + torch.randn_like(mu) * log_std.exp()
                O[t], A[t], V[t] = obs, a, v
                LP[t] = (-0.5 * ((a - mu) / log_std.exp()) ** 2 - log_std - 0.9189385).sum(-1)
                out, rew, done, _ = env.step(a.clamp(-1, 1))
                obs = out["obs"].clone(); R[t] = rew * 0.01; D[t] = done.float()  # reward_shaper scale_value 0.01
            V[horizon] = pi(obs_norm(obs))[2]
            ret_scale = torch.sqrt(ret_norm.var + 1e-5)
            Vd = V * ret_scale  # value head predicts normalised returns
            adv = torch.zeros(num_envs, device=dev); ADV = torch.empty_like(R)
            for t in reversed(range(horizon)):
                nd = 1.0 - D[t]
                delta = R[t] + gamma * Vd[t + 1] * nd - Vd[t]
                adv = delta + gamma * lam * nd * adv
                ADV[t] = adv
            RET = ADV + Vd[:-1]
            ret_norm.update(RET.reshape(-1))
        st = task.read_episode_stats().cpu()
        mean_rew = float(R.mean()) / 0.01
        curve.append((mean_rew, float(st[2] / max(float(st[1]), 1.0))))
        o, a_, lp, adv_, ret_ = O.reshape(-1, n_obs), A.reshape(-1, n_act), LP.reshape(-1), ADV.reshape(-1), RET.reshape(-1)
        adv_ = (adv_ - adv_.mean()) / (adv_.std() + 1e-8)
        ret_n = ret_ / torch.sqrt(ret_norm.var + 1e-5)
        on = obs_norm(o)
        N = o.shape[0]
        for _ in range(epochs):
            perm = torch.randperm(N, device=dev)
            for mb in perm.chunk(minibatches):
                mu, log_std, v = pi(on[mb])
                nlp = (-0.5 * ((a_[mb] - mu) / log_std.exp()) ** 2 - log_std - 0.9189385).sum(-1)
                ratio = (nlp - lp[mb]).exp()
                l_pi = -torch.min(ratio * adv_[mb], ratio.clamp(1 - clip, 1 + clip) * adv_[mb]).mean()
                l_v = 0.5 * (v - ret_n[mb]).pow(2).mean()
                loss = l_pi + 2.0 * l_v + 1e-4 * (mu ** 2).mean()  # bounds loss stand-in
                opt.zero_grad(set_to_none=True); loss.backward()
                nn.utils.clip_grad_norm_(pi.parameters(), 1.0); opt.step()
        if verbose and (it % 10 == 0 or it == iterations - 1):
            print(f"iter {it:4d}  mean reward/step {mean_rew:8.3f}  mean episode length {curve[-1][1]:7.1f}  "
                  f"({(it + 1) * horizon * num_envs / (time.time() - t0) / 1e6:.2f} M env-steps/s incl. learning)", flush=True)
    return curve


if __name__ == "__main__":
    robot = sys.argv[1] if len(sys.argv) > 1 else "Ant"
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 150
    n = int(sys.argv[3]) if len(sys.argv) > 3 else 4096
    c = run(robot, iters, n)
    k = max(len(c) // 10, 1)
    first = sum(x[0] for x in c[:k]) / k; last = sum(x[0] for x in c[-k:]) / k
    print(f"{robot}: mean reward/step first {k} iterations {first:.3f} -> last {k} iterations {last:.3f}")
