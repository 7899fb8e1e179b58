"""TEST INFRASTRUCTURE ONLY -- container-side loader of the VERBATIM reference jit functions.

Imports ``tasks/shared/locomotion.py``, ``tasks/humanoid.py`` and ``tasks/ant.py`` straight
from ``/root/reference`` (read-only, never copied) after installing ``sys.modules`` stubs for
the Isaac Sim packages those files import.  The seven third-party
math helpers come from ``oracle/isaac_math.py``.

``/root/reference`` does not exist on the GPU box, so this module is used ONLY by
``oracle/gen_golden.py`` and by CPU tests that skip when the reference tree is absent.
Nothing in the product package, bench.py or the ``-m gpu`` tests imports it.
"""
import importlib.util
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("OIGE_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "omniisaacgymenvs/tasks/shared/locomotion.py"))


def _mod(name: str, **attrs) -> types.ModuleType:
    m = sys.modules.get(name)
    if m is None:
        m = types.ModuleType(name)
        m.__path__ = []  # behave as a package so sub-imports resolve
        sys.modules[name] = m
    for k, v in attrs.items():
        setattr(m, k, v)
    return m


class _Dummy:
    def __init__(self, *a, **k):
        pass


def install_stubs() -> None:
    from oracle import isaac_math as im

    for pkg in ("omni", "omni.isaac", "omni.isaac.core", "omni.isaac.core.utils", "omni.isaac.core.utils.torch"):
        _mod(pkg)
    _mod(
        "omni.isaac.core.utils.torch.rotations",
        compute_heading_and_up=im.compute_heading_and_up,
        compute_rot=im.compute_rot,
        quat_conjugate=im.quat_conjugate,
    )
    _mod(
        "omni.isaac.core.utils.torch.maths",
        torch_rand_float=im.torch_rand_float,
        tensor_clamp=im.tensor_clamp,
        unscale=im.unscale,
    )
    _mod("omni.isaac.core.articulations", ArticulationView=_Dummy)
    _mod("omni.isaac.core.utils.prims", get_prim_at_path=lambda *a, **k: None)
    _mod("pxr", PhysxSchema=_Dummy)
    # the real RLTask needs gym / omni.kit / cloner / replicator: stub the module it lives in
    for pkg in (
        "omniisaacgymenvs",
        "omniisaacgymenvs.tasks",
        "omniisaacgymenvs.tasks.base",
        "omniisaacgymenvs.tasks.shared",
        "omniisaacgymenvs.robots",
        "omniisaacgymenvs.robots.articulations",
    ):
        _mod(pkg)
    _mod("omniisaacgymenvs.tasks.base.rl_task", RLTask=_Dummy)
    _mod("omniisaacgymenvs.robots.articulations.humanoid", Humanoid=_Dummy)
    _mod("omniisaacgymenvs.robots.articulations.ant", Ant=_Dummy)


def _load(modname: str, relpath: str):
    path = os.path.join(REFERENCE_ROOT, relpath)
    spec = importlib.util.spec_from_file_location(modname, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[modname] = mod
    spec.loader.exec_module(mod)
    return mod


_cache = {}


def load_reference():
    """Returns (locomotion, humanoid, ant) modules executed verbatim from /root/reference."""
    if "mods" not in _cache:
        if not reference_available():
            raise FileNotFoundError(f"reference tree not found at {REFERENCE_ROOT}")
        install_stubs()
        loco = _load("omniisaacgymenvs.tasks.shared.locomotion", "omniisaacgymenvs/tasks/shared/locomotion.py")
        hum = _load("omniisaacgymenvs.tasks.humanoid", "omniisaacgymenvs/tasks/humanoid.py")
        ant = _load("omniisaacgymenvs.tasks.ant", "omniisaacgymenvs/tasks/ant.py")
        _cache["mods"] = (loco, hum, ant)
    return _cache["mods"]
