"""Re-captures the ncu summary bench.py quotes for the dominant kernel, from the binary that is in the tree NOW.

Run it as the LAST GPU call of a round (on the GPU box, from the repo root):

    python tools/recapture_k2f.py            # writes profiles/r3_ncu_k2f_sc_16384.txt

It runs `ncu --set full --clock-control none -k regex:env_step_role -s 20 -c 1 python tools/k2f_once.py Humanoid 16384`
(K2-F, 16384 Humanoids, self-collision on: the launch bench.py times), keeps the metrics the design discussion uses
and heads the file with the SHA-256 of the kernel sources (`kernel_source_hash()`; `.git` does not travel to the GPU
box, so a source hash stands in for the commit).  bench.py compares that hash with the sources it runs from and labels
the figures `"source": "stale"` when they differ.  Numbers printed under ncu are never bench values: only the
per-launch counters (instructions, DRAM bytes, stall ratios) are taken from here.
"""
import csv
import hashlib
import io
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles", "r3_ncu_k2f_sc_16384.txt")
KEEP_PREFIXES = ("dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum", "launch__", "lts__t_sector_hit_rate.pct",
                 "l1tex__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
                 "sm__warps_active.avg.per_cycle_active", "smsp__average_warps_issue_stalled_", "smsp__cycles_active.avg",
                 "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
                 "smsp__sass_inst_executed_op_local", "smsp__sass_inst_executed_op_shared", "sm__throughput.avg.pct_of_peak_sustained_elapsed")


def kernel_source_hash() -> str:
    """SHA-256 over the CUDA sources and the C header, in name order (what decides the machine code of the kernels)."""
    h = hashlib.sha256()
    csrc = os.path.join(ROOT, "omniisaacgymenvs_b200", "csrc")
    files = sorted(os.path.join(csrc, f) for f in os.listdir(csrc) if f.endswith((".cu", ".cuh")))
    files.append(os.path.join(ROOT, "include", "oige_b200.h"))
    for p in files:
        h.update(os.path.basename(p).encode())
        h.update(open(p, "rb").read())
    return h.hexdigest()[:16]


def main() -> int:
    cmd = ["ncu", "--set", "full", "--clock-control", "none", "-k", "regex:env_step_role", "-s", "20", "-c", "1",
           "--csv", "--page", "raw", sys.executable, os.path.join(ROOT, "tools", "k2f_once.py"), "Humanoid", "16384"]
    r = subprocess.run(cmd, capture_output=True, text=True, cwd=ROOT)
    if r.returncode != 0:
        sys.stderr.write(r.stdout[-2000:] + r.stderr[-2000:])
        return r.returncode
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith('"')]
    rows = list(csv.reader(io.StringIO("\n".join(lines))))
    if len(rows) < 3:
        sys.stderr.write("no ncu CSV rows found\n" + r.stdout[-2000:])
        return 1
    names, units, vals = rows[0], rows[1], rows[2]
    kern = vals[names.index("Kernel Name")] if "Kernel Name" in names else "env_step_role_kernel"
    out = [f"# source_hash {kernel_source_hash()}",
           "# " + " ".join(cmd[:11]) + " python tools/k2f_once.py Humanoid 16384   (tools/recapture_k2f.py, round 3)",
           f"# {kern}: the WHOLE env step of 16384 Humanoid envs (self-collision on) in one launch"]
    for nm, un, v in sorted(zip(names, units, vals)):
        if nm == "gpu__time_duration.sum" and un == "ns":  # bench.py and the earlier summaries quote microseconds
            v, un = f"{float(v.replace(',', '')) / 1e3:.3f}", "us"
        if nm.startswith(KEEP_PREFIXES):
            out.append(f"{nm:<85} {v} {un}".rstrip())
    with open(OUT, "w") as f:
        f.write("\n".join(out) + "\n")
    gp = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(gp):
        with open(os.path.join(gp, os.path.basename(OUT)), "w") as f:
            f.write("\n".join(out) + "\n")
    print("\n".join(out[:3]))
    print(f"wrote {OUT} ({len(out) - 3} metrics)")
    return 0


if __name__ == "__main__":
    sys.exit(main())
