"""Diagnostic (GPU box): K2 vs float64 oracle error statistics."""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import dyn_helpers as dh
from omniisaacgymenvs_b200 import _lib
lib = _lib.load()
for robot in ("Humanoid", "Ant"):
    model, spec = dh.specs(robot)
    for label, h, calls in (("free 1 substep", 5.0, None), ("contact 1 substep", 1.25 if robot == "Humanoid" else 0.3, None), ("free 64", 50.0, 32), ("contact 64", model.spawn_height, 32)):
        st = dh.random_state(model, 1024, 0, height=h, tilt=0.15)
        if calls is None:
            g = dh.run_gpu(lib, model, spec, st, 1, 0.0083); o = dh.run_oracle(model, spec, st, 1, 0.0083)
        else:
            g = dh.run_gpu(lib, model, spec, st, 2, 0.0083, calls=calls); o = dh.run_oracle(model, spec, st, 2, 0.0083, calls=calls)
        out = [robot, label]
        for k in ("q", "qd", "pos", "vel"):
            e = np.abs(g[k] - o[k])
            out.append(f"{k}: max {e.max():.2e} med {np.median(e):.2e} p99 {np.quantile(e, 0.99):.2e}")
        # acceleration-level relative error for single step: (qd1 - qd0)
        if calls is None:
            dq_g = g["qd"] - st["qd"]; dq_o = o["qd"] - st["qd"]
            rel = np.abs(dq_g - dq_o) / np.maximum(np.abs(dq_o), 1e-3)
            out.append(f"dqd rel: max {rel.max():.2e} med {np.median(rel):.2e}")
        print(" | ".join(out))
