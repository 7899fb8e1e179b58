"""TEST / MEASUREMENT INFRASTRUCTURE ONLY (like the rest of oracle/) -- CPU baseline timings of SURVEY.md section 8(d) / BASELINE.json configs[0] (run in the BUILD container, where
/root/reference exists -- nothing here runs on the GPU box):
  (i)  the VERBATIM reference torch.jit functions get_observations + get_dof_at_limit_cost + calculate_metrics + is_done
       (tasks/shared/locomotion.py:194-321, tasks/humanoid.py:120-127, tasks/ant.py:92-95), Ant and Humanoid at 4096 envs,
       torch.set_num_threads(1) and all cores, >= 50 iterations after 3 warm-ups;
  (ii) the float64 dynamics oracle (oracle/dynamics_oracle.c) per env-substep, 1 thread and all cores.
Output: profiles/r2_cpu_reference_timing.txt (stdout).  Run as `python -m oracle.cpu_ref_time` from the repo root."""
import os, sys, time, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)  # repo root
import numpy as np
import torch
from oracle import locomotion_oracle as lo, ref_loader, dynamics as dyn
from oracle.gen_golden import run_reference
from omniisaacgymenvs_b200.robots.model import build_model

assert ref_loader.reference_available(), "needs /root/reference (build container)"
ncores = os.cpu_count() or 1
print(f"# host: {ncores} logical cores, torch {torch.__version__}, {time.strftime('%Y-%m-%d')}")
print("# (i) verbatim reference jit functions, one post_physics_step pass (obs + limit cost + reward + done)")
for robot in ("Ant", "Humanoid"):
    spec = lo.SPECS[robot]
    for n in (4096,):
        s = lo.synthetic_state(spec, n, seed=0)
        for threads in (1, ncores):
            torch.set_num_threads(threads)
            for _ in range(3): run_reference(spec, s)
            ts = []
            for _ in range(60):
                t0 = time.perf_counter(); run_reference(spec, s); ts.append(time.perf_counter() - t0)
            med = statistics.median(ts)
            print(f"{robot:9s} {n:6d} envs  threads={threads:3d}  median {med*1e3:8.3f} ms  min {min(ts)*1e3:8.3f} ms  "
                  f"{n/med/1e6:7.3f} M env-steps/s  ({60} iterations after 3 warm-ups)")
print("# (ii) float64 dynamics oracle (dense projection method, incl. self-collision for the Humanoid), per env-substep")
for robot in ("Ant", "Humanoid"):
    m = build_model(robot); desc = dyn.make_desc(m); D = m.num_dof
    n = 4096
    rng = np.random.default_rng(0); lim = m.dof_limits().astype(np.float64)
    for threads in (1, ncores):
        pos = np.zeros((n, 3)); pos[:, 2] = m.spawn_height
        quat = np.tile([[1.0, 0, 0, 0]], (n, 1)); vel = np.zeros((n, 6))
        q = np.clip(rng.uniform(-0.2, 0.2, (n, D)), lim[:, 0], lim[:, 1]); qd = rng.uniform(-0.1, 0.1, (n, D))
        eff = np.ascontiguousarray(rng.uniform(-1, 1, (n, D)) * 10.0); sens = np.zeros((n, m.num_sensors, 6))
        dyn.step(desc, pos, quat, vel, q, qd, eff, sens, 2, 1 / 120, threads)
        reps = 3 if threads == 1 else 20
        t0 = time.perf_counter()
        for _ in range(reps): dyn.step(desc, pos, quat, vel, q, qd, eff, sens, 2, 1 / 120, threads)
        dt = (time.perf_counter() - t0) / reps
        print(f"{robot:9s} {n:6d} envs  threads={threads:3d}  {dt*1e3:9.2f} ms per 2-substep call  {dt/(2*n)*1e6:8.3f} us per env-substep"
              f"  {n/dt/1e6:7.4f} M env-steps/s")
