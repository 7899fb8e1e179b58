"""K2 kernel variants side by side (GPU box): device time of the articulation step (2 substeps) per mode and env count."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
robots = sys.argv[1] if len(sys.argv) > 1 else "Humanoid,Ant"
ns = sys.argv[2] if len(sys.argv) > 2 else "1024,2048,4096,8192,16384,32768,65536"
for mode in os.environ.get("K2_MODES", "warp,thread,role").split(","):
    env = dict(os.environ, OIGE_K2_MODE=mode, K2_NS=ns)
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "k2_time.py"), robots], env=env, capture_output=True, text=True)
    for line in out.stdout.splitlines():
        print(f"[{mode:6s}] {line}", flush=True)
    if out.returncode != 0:
        print(f"[{mode}] FAILED\n{out.stderr[-2000:]}", flush=True)
