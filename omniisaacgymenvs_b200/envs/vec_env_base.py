"""Headless stand-ins for the Isaac Sim runtime objects the reference's env wrapper drives:
``omni.isaac.gym.vec_env.VecEnvBase``, ``omni.isaac.core.World`` and ``SimulationApp``
(used by envs/vec_env_rlgames.py:30-89 and scripts/random_policy.py:53-63).

No Kit, no USD stage, no rendering: ``World.step`` advances the articulation kernel by one
``sim.dt`` and nothing else.
"""
import copy
import ctypes as C
from typing import Optional

import numpy as np
import torch

from omniisaacgymenvs_b200 import _lib


class _HeadlessApp:
    """``env._simulation_app`` for scripts that poll it (scripts/random_policy.py:53)."""

    def is_running(self) -> bool:
        return True

    def close(self) -> None:
        return None


class Scene:
    """Minimal ``omni.isaac.core.scenes.Scene``: remembers the views tasks add to it."""

    def __init__(self):
        self._objects = {}

    def add(self, obj):
        self._objects[getattr(obj, "name", str(len(self._objects)))] = obj
        return obj

    def add_default_ground_plane(self, prim_path: str = "/World/defaultGroundPlane", **kw):
        self._objects["ground_plane"] = prim_path
        return prim_path

    def get_object(self, name):
        return self._objects.get(name)


class World:
    """Steps the physics of the task's articulation view (stand-in for ``omni.isaac.core.World``)."""

    def __init__(self, physics_dt: float, device: str, sim_params: Optional[dict] = None):
        self._physics_dt = float(physics_dt)
        self._device = device
        self._sim_params = sim_params or {}
        self.scene = Scene()
        self._task = None
        self.current_time_step_index = 0
        self._playing = True

    def add_task(self, task) -> None:
        self._task = task

    def get_physics_dt(self) -> float:
        return self._physics_dt

    def is_playing(self) -> bool:
        return self._playing

    def stop(self) -> None:
        self._playing = False

    def play(self) -> None:
        self._playing = True

    def reset(self, soft: bool = False) -> None:
        """Builds the scene (first call) and runs ``task.post_reset`` (World.reset semantics)."""
        if self._task is None:
            return
        if not soft or getattr(self._task, "_robots", None) is None:
            self._task.set_up_scene(self.scene)
        self._task.post_reset()
        if hasattr(self._task, "set_up_domain_randomization"):
            self._task.set_up_domain_randomization()
        self.current_time_step_index = 0

    def step(self, render: bool = False) -> None:
        """One physics step of ``sim.dt`` with the joint efforts last written to the view."""
        view = self._task.get_robot()
        lib = _lib.load()
        rc = lib.oige_articulation_step_f32(
            view.model_handle, C.byref(self._task._unit_gear_params), C.byref(view.state_struct),
            view.dof_efforts.data_ptr(), view.count, 1, self._physics_dt, _lib.current_stream_ptr(view._device))
        _lib.check(rc, "oige_articulation_step_f32")
        self.current_time_step_index += 1


class Box:
    """Tiny stand-in for ``gym.spaces.Box`` (gym is not installed): what rl_games reads."""

    def __init__(self, low, high):
        self.low = np.asarray(low, dtype=np.float32)
        self.high = np.asarray(high, dtype=np.float32)
        self.shape = self.low.shape
        self.dtype = np.float32

    def sample(self):
        lo = np.where(np.isfinite(self.low), self.low, -1.0)
        hi = np.where(np.isfinite(self.high), self.high, 1.0)
        return np.random.uniform(lo, hi).astype(np.float32)

    def __repr__(self):
        return f"Box({self.low.min()}, {self.high.max()}, {self.shape}, float32)"


class VecEnvBase:
    """Stand-in for ``omni.isaac.gym.vec_env.VecEnvBase`` (headless only)."""

    def __init__(self, headless: bool = True, sim_device: int = 0, **kwargs):
        self._simulation_app = _HeadlessApp()
        self._render = False  # no graphics stack on this platform; `headless` is accepted and ignored
        self._headless = headless
        self._sim_device = sim_device
        self.sim_frame_count = 0
        self._world = None
        self._task = None

    def set_task(self, task, backend="numpy", sim_params=None, init_sim=True) -> None:
        if backend != "torch":
            raise RuntimeError("only backend='torch' is supported (utils/task_util.py:70 passes 'torch')")
        sim_params = copy.deepcopy(sim_params) if sim_params else {}
        device = sim_params.get("sim_device", f"cuda:{self._sim_device}")
        self._world = World(physics_dt=sim_params.get("dt", 1.0 / 60.0), device=device, sim_params=sim_params)
        self._world.add_task(task)
        self._task = task
        self._num_envs = task.num_envs
        self.observation_space = task.observation_space
        self.action_space = task.action_space
        if init_sim:
            self._world.reset()

    @property
    def num_envs(self) -> int:
        return self._num_envs

    def get_number_of_agents(self) -> int:  # utils/rlgames/rlgames_utils.py:104-105 calls it
        return getattr(self._task, "num_agents", 1)

    def render(self, mode="human") -> None:
        return None

    def close(self) -> None:
        self._simulation_app.close()

    def seed(self, seed=-1):
        from omniisaacgymenvs_b200.utils.maths import set_seed

        return set_seed(seed)
