"""Measured deviation of the fused obs / reward / termination kernel (K1, and the same arithmetic in the whole-step
kernel's epilogue) from the oracle -- which is itself bit-identical to the verbatim reference functions
(tests/test_oracle_golden.py).  Reference arithmetic: tasks/shared/locomotion.py:194-321, tasks/humanoid.py:120-127,
tasks/ant.py:92-95.

`python tests/k1_deviation.py [out.txt]` (needs a B200) prints, per robot, over 4096 envs x 4 seeds and over the golden
fixtures: for every output the number of elements that are not bit-identical, the largest |delta| and the largest
relative deviation.  The table is the source of the constants in helpers.assert_k1_matches and of DESIGN.md section 4.
"""
import math
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import helpers  # noqa: E402
from oracle import locomotion_oracle as lo  # noqa: E402

ANGLE_COLS = (7, 8, 9)


def columns(spec):
    d, s6 = spec.num_dof, 6 * spec.num_sensors
    return {
        "obs[0] height": [0],
        "obs[1:7] local velocities": list(range(1, 7)),
        "obs[7:10] yaw/roll/angle_to_target (circle)": list(ANGLE_COLS),
        "obs[10] up_proj": [10],
        "obs[11] heading_proj": [11],
        "obs dof_pos_scaled": list(range(12, 12 + d)),
        "obs dof_vel": list(range(12 + d, 12 + 2 * d)),
        "obs sensors": list(range(12 + 2 * d, 12 + 2 * d + s6)),
        "obs actions": list(range(12 + 2 * d + s6, 12 + 3 * d + s6)),
    }


class Acc:
    def __init__(self):
        self.n = 0
        self.nbits = 0
        self.max_abs = 0.0
        self.max_rel = 0.0

    def add(self, got, ref, circle=False):
        g, r = got.double().flatten(), ref.double().flatten()
        d = (g - r).abs()
        if circle:
            d = torch.minimum(d, (2 * math.pi - d).abs())
        self.n += g.numel()
        self.nbits += int((got.flatten() != ref.flatten()).sum())
        if g.numel():
            self.max_abs = max(self.max_abs, float(d.max()))
            # relative to |ref| with the 1e-5 absolute floor of the parity bar's O(1) quantities
            self.max_rel = max(self.max_rel, float((d / r.abs().clamp_min(1.0)).max()))


def measure(cases, device="cuda:0", closed_form=1):
    """cases: iterable of (spec, state dict, reference outputs).  Returns {output name: Acc} per robot."""
    table = {}
    for spec, s, ref in cases:
        got = helpers.run_k1(spec, s, device, closed_form_angles=closed_form)
        t = table.setdefault(spec.name, {})
        for name, cols in columns(spec).items():
            t.setdefault(name, Acc()).add(got["obs"][:, cols], ref["obs"][:, cols], circle="circle" in name)
        for k in ("rew", "potentials", "prev_potentials", "up_vec", "heading_vec"):
            t.setdefault(k, Acc()).add(got[k], ref[k])
        for k in ("reset_buf", "progress_buf"):
            t.setdefault(k, Acc()).add(got[k].float(), ref[k].float())
        # reward deviation on the rows whose potentials ARE bit-identical (separates the reward's own arithmetic from
        # the cancellation of two ~6e4 potentials)
        same = got["potentials"] == ref["potentials"]
        t.setdefault("rew | potentials bit-equal", Acc()).add(got["rew"][same], ref["rew"][same])
    return table


def synthetic_cases(n=4096, seeds=(0, 1, 2, 3)):
    for robot in ("Humanoid", "Ant"):
        spec = lo.SPECS[robot]
        for seed in seeds:
            s = lo.synthetic_state(spec, n, seed)
            yield spec, s, lo.post_physics_step(spec, {k: v.clone() for k, v in s.items()})


def golden_cases():
    for name in helpers.golden_names():
        s, ref = helpers.load_golden(name)
        yield lo.SPECS["Humanoid" if name.startswith("humanoid") else "Ant"], s, ref


def render(title, table):
    lines = [title]
    for robot, t in table.items():
        lines.append(f"  {robot}")
        lines.append(f"    {'output':46s} {'elements':>9s} {'not bit-equal':>13s} {'max |d|':>11s} {'max rel (floor 1)':>18s}")
        for name, a in t.items():
            lines.append(f"    {name:46s} {a.n:9d} {a.nbits:13d} {a.max_abs:11.3e} {a.max_rel:18.3e}")
    return "\n".join(lines)


def main():
    out = []
    for mode, label in ((1, "closed-form angle wrap (default)"), (0, "literal atan2(sin, cos) chain")):
        out.append(render(f"== synthetic 4096 envs x 4 seeds, {label}", measure(synthetic_cases(), closed_form=mode)))
        out.append(render(f"== golden fixtures (tests/golden), {label}", measure(golden_cases(), closed_form=mode)))
    text = "\n".join(out)
    print(text)
    if len(sys.argv) > 1:
        with open(sys.argv[1], "w") as f:
            f.write(text + "\n")


if __name__ == "__main__":
    main()
