"""Builds ``liboige_b200.so`` (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

    python -m omniisaacgymenvs_b200.build [--force] [--verbose]

nvcc cross-compiles without a GPU.  The resulting ``.so`` is git-ignored but travels to the
GPU box with the repo snapshot.  K1/K3 are compiled with ``-fmad=false`` so their arithmetic
follows the reference's operation order (see csrc/obs_reward_reset.cu); K2 keeps FMA contraction.
"""
import argparse
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
REPO_ROOT = os.path.dirname(PKG_DIR)
CSRC = os.path.join(PKG_DIR, "csrc")
INCLUDE = os.path.join(REPO_ROOT, "include")
OBJ_DIR = os.path.join(CSRC, "build")
LIB_PATH = os.path.join(PKG_DIR, "liboige_b200.so")

ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-I", INCLUDE, "-I", CSRC]
# (source, extra flags)
SOURCES = [
    ("obs_reward_reset.cu", ["-fmad=false"]),
    ("reset_idx.cu", ["-fmad=false"]),
    ("cartpole.cu", ["-fmad=false"]),
    ("anymal.cu", ["-fmad=false"]),
    ("domain_randomization.cu", ["-fmad=false"]),
    ("articulation_step.cu", []),
    ("capi.cu", []),
    ("env_step_host_static.cu", []),
]


def _nvcc() -> str:
    cand = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(cand):
        raise RuntimeError("nvcc not found: the CUDA toolkit is required to build liboige_b200.so")
    return cand


def _deps():
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hdrs.append(os.path.join(INCLUDE, "oige_b200.h"))
    hdrs.append(os.path.abspath(__file__))
    return hdrs


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build_library(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ_DIR, exist_ok=True)
    nvcc = _nvcc()
    hdrs = _deps()
    jobs = []
    objs = []
    for src, extra in SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ_DIR, src.replace(".cu", ".o"))
        objs.append(o)
        if force or _stale(o, [s] + hdrs):
            cmd = [nvcc] + ARCH_FLAGS + COMMON + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", s, "-o", o]
            jobs.append(cmd)

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
        return r.stderr

    if jobs:
        with ThreadPoolExecutor(max_workers=len(jobs)) as ex:
            for out in ex.map(run, jobs):
                if verbose and out:
                    print(out)
    if jobs or force or _stale(LIB_PATH, objs):
        cmd = [nvcc] + ARCH_FLAGS + ["-shared", "-o", LIB_PATH] + objs
        run(cmd)
    return LIB_PATH


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args()
    print(build_library(force=a.force, verbose=a.verbose))
    sys.exit(0)
