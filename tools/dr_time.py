"""Observation / action noise kernel alone at bandwidth sizes (GPU box): `oige_dr_apply_f32`, both blocks active
(on_reset gaussian + on_interval gaussian every step, in-kernel Philox), ~3 % of the envs flagged for reset.
Algorithmic bytes per env-step: buffer read + written 8*dim, correlated noise read 4*dim, flag 8, counter 4 + 4.

    python tools/dr_time.py > gpurun_out/dr_time.txt
"""
import ctypes as C, json, math, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np
import torch
from omniisaacgymenvs_b200 import _lib

lib = _lib.load(); dev = torch.device("cuda:0")
try:
    peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    peak = 6551.0
flush = torch.zeros(64 << 20, device=dev)
print(f"# {torch.cuda.get_device_name(0)}; HBM peak {peak:.0f} GB/s; dr_apply_kernel, L2 flushed, median of 30")
for dist, dname in ((_lib.DR_GAUSSIAN, "gaussian"), (_lib.DR_UNIFORM, "uniform")):
    for dim in (87, 60, 21):
        for n in (16384, 1 << 20, 1 << 22):
            r = _lib.OigeNoiseSpec(1, _lib.DR_ADDITIVE, dist, 1, 0.0, 0.01)
            i = _lib.OigeNoiseSpec(1, _lib.DR_ADDITIVE, dist, 1, 0.0, 0.05)
            cnt = torch.zeros(n, dtype=torch.int32, device=dev); corr = torch.zeros(n, dim, device=dev)
            st = _lib.OigeNoiseState(cnt.data_ptr(), corr.data_ptr())
            buf = torch.randn(n, dim, device=dev); rb = (torch.rand(n, device=dev) < 0.03).long()
            ts = []
            for it in range(35):
                flush.add_(1.0)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                rc = lib.oige_dr_apply_f32(C.byref(r), C.byref(i), C.byref(st), buf.data_ptr(), buf.data_ptr(), rb.data_ptr(), n, dim,
                                           math.inf, 5.0, None, None, 7, it + 1, None, _lib.DR_STREAM_OBSERVATIONS, 0,
                                           _lib.current_stream_ptr(dev))
                e1.record(); torch.cuda.synchronize(); _lib.check(rc, "dr")
                if it >= 5: ts.append(e0.elapsed_time(e1))
            ms = float(np.median(ts)); b = 12 * dim + 16
            print(f"{dname:8s} dim={dim:3d} n={n:8d}  {ms*1e3:8.1f} us  {b*n/ms/1e6:7.0f} GB/s  frac {b*n/ms/1e6/peak:.2f}", flush=True)
