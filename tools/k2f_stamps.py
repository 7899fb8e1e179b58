"""Per-role phase timeline of K2-F (env_step_role_kernel) from clock64 stamps of group 0 of CTA 0.
Build (here, nvcc cross-compiles):  python tools/k2f_stamps.py build   -> tools/_stamps/liboige_b200_stamps.so
Run (GPU box):                      python tools/k2f_stamps.py [robot] [envs]
The diagnostic library is the product library compiled with -DOIGE_STAMPS; it is never loaded by the package itself."""
import ctypes as C, os, subprocess, sys, io, contextlib
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "tools", "_stamps"); LIB = os.path.join(OUT, "liboige_b200_stamps.so")

def build():
    from omniisaacgymenvs_b200 import build as B
    os.makedirs(OUT, exist_ok=True)
    objs = []
    for src, extra in B.SOURCES:
        o = os.path.join(OUT, src.replace(".cu", ".o")); objs.append(o)
        subprocess.run([B._nvcc()] + B.ARCH_FLAGS + B.COMMON + extra + ["-DOIGE_STAMPS", "-c", os.path.join(B.CSRC, src), "-o", o], check=True)
    subprocess.run([B._nvcc()] + B.ARCH_FLAGS + ["-shared", "-o", LIB] + objs, check=True)
    print(LIB)

TAGS = {1: "kernel: state in registers", 10: "substep start", 40: "pairs: broad done", 41: "pairs: narrow done", 42: "pairs: hit barrier passed",
        43: "pairs: contacts done", 44: "pairs: barrier passed", 80: "S2b done / at barrier A", 81: "barrier A passed", 90: "substeps done", 91: "all groups done (syncthreads)", 92: "output tile written", 93: "obs quarters done", 94: "reward / stats done", 95: "fence + sync before the bulk stores", 99: "bulk stores issued"}
def name(t):
    if t in TAGS: return TAGS[t]
    k, sg = divmod(t, 10)
    return {2: f"seg {sg}: start", 3: f"seg {sg}: barrier-before passed", 5: f"seg {sg}: kinematics done", 6: f"seg {sg}: sweep done",
            7: f"seg {sg}: barrier-after passed"}.get(k, str(t))

if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "build":
        build(); sys.exit(0)
    import torch
    from omniisaacgymenvs_b200 import _lib
    _lib.LIB_PATH = LIB
    from omniisaacgymenvs_b200.utils.task_util import make_env
    robot = sys.argv[1] if len(sys.argv) > 1 else "Humanoid"
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
    lib = _lib.load()
    with contextlib.redirect_stdout(io.StringIO()):
        env, task = make_env(robot, num_envs=n); env.reset()
    dev = torch.device(task.device)
    for _ in range(30):
        env.step(torch.rand(n, task.num_actions, device=dev) * 2 - 1)
    torch.cuda.synchronize()
    lib.oige_debug_clear_stamps()
    env.step(torch.rand(n, task.num_actions, device=dev) * 2 - 1)
    torch.cuda.synchronize()
    buf = (C.c_longlong * (4 * 128 * 2))()
    assert lib.oige_debug_read_stamps(buf) == 0
    import numpy as np
    st = np.array(buf[:]).reshape(4, 128, 2)
    t0 = min(st[r, 0, 1] for r in range(4) if st[r, 0, 0] != 0)
    for r in range(4):
        if st[r, 0, 0] == 0: continue
        print(f"--- role {r} (cycles since the first role's first stamp; delta to the previous stamp)")
        prev = t0
        for tag, clk in st[r]:
            if tag == 0: break
            print(f"  {clk - t0:8d} (+{clk - prev:6d})  {name(int(tag))}")
            prev = clk
