"""TEST INFRASTRUCTURE ONLY -- Python face of the float64 CPU dynamics oracle (oracle/dynamics_oracle.c).

Compiles the C file with gcc into ``oracle/_build/liboracle_dyn.so`` on first use and exposes
``step()`` on numpy float64 arrays.  See the C file header for the substitution notice (PhysX
is unavailable; this integrator is the documented stand-in oracle for K2).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
_SRC = os.path.join(_HERE, "dynamics_oracle.c")
_BUILD = os.path.join(_HERE, "_build")
_LIB = os.path.join(_BUILD, "liboracle_dyn.so")
_lib = None


def build(force: bool = False) -> str:
    os.makedirs(_BUILD, exist_ok=True)
    hdr = os.path.join(_ROOT, "include", "oige_b200.h")
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < max(os.path.getmtime(_SRC), os.path.getmtime(hdr)):
        cmd = ["gcc", "-O2", "-fopenmp", "-shared", "-fPIC", "-I", os.path.join(_ROOT, "include"), _SRC, "-o", _LIB, "-lm"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("gcc failed building the dynamics oracle:\n" + r.stdout + r.stderr)
    return _LIB


def _load():
    global _lib
    if _lib is None:
        lib = C.CDLL(build())
        dp = C.POINTER(C.c_double)
        lib.oracle_articulation_step_f64.argtypes = [C.c_void_p, C.c_int64, dp, dp, dp, dp, dp, dp, dp, dp, C.c_int,
                                                     C.c_double, C.c_int]
        lib.oracle_articulation_step_f64.restype = C.c_int
        lib.oracle_energy_f64.argtypes = [C.c_void_p, dp, dp, dp, dp, dp]
        lib.oracle_energy_f64.restype = C.c_double
        _lib = lib
    return _lib


def _p(a):
    return a.ctypes.data_as(C.POINTER(C.c_double)) if a is not None else None


def make_desc(model):
    """OigeModelDesc (ctypes) for an ArticulationModel -- the struct layout is the public ABI's."""
    from omniisaacgymenvs_b200 import _lib as abi  # struct definitions only; no CUDA needed
    from omniisaacgymenvs_b200.robots.model import fill_desc

    desc = abi.OigeModelDesc()
    fill_desc(model, desc)
    return desc


def step(desc, pos, quat, vel, q, qd, efforts, sensors=None, substeps=1, dt=0.0083, num_threads=1, targets=None) -> int:
    """In-place step of float64 C-contiguous arrays pos[N,3] quat[N,4] vel[N,6] q[N,D] qd[N,D]."""
    for a in (pos, quat, vel, q, qd, efforts):
        assert a.dtype == np.float64 and a.flags.c_contiguous
    n = pos.shape[0]
    return _load().oracle_articulation_step_f64(C.byref(desc), n, _p(pos), _p(quat), _p(vel), _p(q), _p(qd),
                                                _p(efforts), _p(targets), _p(sensors), int(substeps), float(dt),
                                                int(num_threads))


def energy(desc, pos, quat, vel, q, qd) -> float:
    return _load().oracle_energy_f64(C.byref(desc), _p(pos), _p(quat), _p(vel), _p(q), _p(qd))
