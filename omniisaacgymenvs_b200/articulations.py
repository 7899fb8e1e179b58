"""ArticulationView-style state container backed by the K2 articulation kernel.

Mirrors the subset of ``omni.isaac.core.articulations.ArticulationView`` the locomotion tasks call
(call sites tasks/shared/locomotion.py:81-89,111-114,130-134,149-150,
tasks/humanoid.py:110-112).  State lives in field-major ``[N,k]`` row-major fp32 tensors -- the
layout the reference's getters hand out -- so each env tile of each field is one contiguous chunk
for the kernels' bulk copies, and ``clone=False`` getters return the live buffers with no copy.
"""
import ctypes as C
from typing import Optional, Tuple

import numpy as np
import torch

from omniisaacgymenvs_b200 import _lib
from omniisaacgymenvs_b200.robots.model import ArticulationModel, fill_desc


def grid_env_origins(num_envs: int, spacing: float, offset: int = 0, count: Optional[int] = None) -> np.ndarray:
    """Env origins laid out like ``omni.isaac.cloner.GridCloner`` (tasks/base/rl_task.py:92-94,
    124-128): near-square grid centred on the world origin.  ``offset`` /
    ``count`` select the rows of one shard so per-env values do not depend on the GPU count."""
    num_per_row = max(int(np.sqrt(num_envs)), 1)
    num_rows = int(np.ceil(num_envs / num_per_row))
    num_cols = int(np.ceil(num_envs / num_rows))
    row_offset = 0.5 * spacing * (num_rows - 1)
    col_offset = 0.5 * spacing * (num_cols - 1)
    idx = np.arange(offset, offset + (num_envs - offset if count is None else count))
    row, col = idx // num_cols, idx % num_cols
    out = np.zeros((len(idx), 3), np.float32)
    out[:, 0] = row_offset - row * spacing
    out[:, 1] = col * spacing - col_offset
    return out


class _PhysicsView:
    """Stand-in for ``ArticulationView._physics_view`` (locomotion.py:89)."""

    def __init__(self, view: "ArticulationView"):
        self._view = view

    def get_force_sensor_forces(self) -> torch.Tensor:
        return self._view.sensor_forces


class ArticulationView:
    def __init__(self, model: ArticulationModel, num_envs: int, device, env_origins: torch.Tensor,
                 name: str = "articulation_view"):
        self.name = name
        self.model = model
        self.count = int(num_envs)
        self.num_dof = model.num_dof
        self._device = torch.device(device)
        if self._device.type != "cuda":
            raise RuntimeError("ArticulationView requires a CUDA device (B200 build, no CPU fallback)")
        lib = _lib.load()
        desc = _lib.OigeModelDesc()
        fill_desc(model, desc)
        with torch.cuda.device(self._device):
            handle = lib.oige_model_create(C.byref(desc))
        if not handle:
            raise RuntimeError("oige_model_create failed: " + lib.oige_last_error().decode())
        self._model_handle = handle
        n, d, s = self.count, self.num_dof, model.num_sensors
        f = dict(device=self._device, dtype=torch.float32)
        self.root_pos = torch.zeros((n, 3), **f)
        self.root_quat = torch.zeros((n, 4), **f)
        self.root_vel = torch.zeros((n, 6), **f)
        self.dof_pos = torch.zeros((n, d), **f)
        self.dof_vel = torch.zeros((n, d), **f)
        self.dof_efforts = torch.zeros((n, d), **f)
        self.sensor_forces = torch.zeros((n, s, 6), **f)
        # default (spawn) state: env origin + spawn height, identity rotation, zero joints
        self.root_pos.copy_(env_origins.to(self._device))
        self.root_pos[:, 2] += model.spawn_height
        self.root_quat[:, 0] = 1.0
        self._dof_limits = torch.from_numpy(model.dof_limits()).to(self._device)
        self._physics_view = _PhysicsView(self)
        self.dof_pos_targets = None  # allocated by the first set_joint_position_targets (PD position drive)
        self._state_struct = _lib.OigeState(
            self.root_pos.data_ptr(), self.root_quat.data_ptr(), self.root_vel.data_ptr(), self.dof_pos.data_ptr(),
            self.dof_vel.data_ptr(), self.sensor_forces.data_ptr(), None)

    def __del__(self):
        h = getattr(self, "_model_handle", None)
        if h:
            try:
                _lib.load().oige_model_destroy(h)
            except Exception:
                pass
            self._model_handle = None

    # ---- C-ABI views ----
    @property
    def state_struct(self) -> "_lib.OigeState":
        return self._state_struct

    @property
    def model_handle(self) -> int:
        return self._model_handle

    # ---- getters (clone=False returns the live buffer, as the physics tensor API does) ----
    def get_world_poses(self, indices=None, clone: bool = True) -> Tuple[torch.Tensor, torch.Tensor]:
        p, q = self.root_pos, self.root_quat
        if indices is not None:
            return p[indices], q[indices]
        return (p.clone(), q.clone()) if clone else (p, q)

    def get_velocities(self, indices=None, clone: bool = True) -> torch.Tensor:
        v = self.root_vel
        if indices is not None:
            return v[indices]
        return v.clone() if clone else v

    def get_joint_positions(self, indices=None, clone: bool = True) -> torch.Tensor:
        x = self.dof_pos
        if indices is not None:
            return x[indices]
        return x.clone() if clone else x

    def get_joint_velocities(self, indices=None, clone: bool = True) -> torch.Tensor:
        x = self.dof_vel
        if indices is not None:
            return x[indices]
        return x.clone() if clone else x

    def get_dof_index(self, dof_name: str) -> int:
        """Column of a named joint in the [N,D] tensors (tasks/cartpole.py:133-134)."""
        return self.model.dof_names.index(dof_name)

    def get_dof_limits(self) -> torch.Tensor:
        return self._dof_limits.unsqueeze(0).expand(self.count, -1, -1)

    # ---- setters (indexed scatter, like the tensor API) ----
    @staticmethod
    def _idx(indices):
        return slice(None) if indices is None else indices.long()

    def set_joint_efforts(self, efforts: torch.Tensor, indices=None) -> None:
        self.dof_efforts[self._idx(indices)] = efforts

    def set_joint_position_targets(self, positions: torch.Tensor, indices=None) -> None:
        """PD position targets (the Anymal-style actuation, tasks/anymal.py:185-187): from the first call on the
        articulation kernel applies tau = kp (target - q) - kd qd with the model's drive gains, on top of efforts."""
        if self.dof_pos_targets is None:
            self.dof_pos_targets = self.dof_pos.clone()
            self._state_struct.dof_pos_targets = self.dof_pos_targets.data_ptr()
        self.dof_pos_targets[self._idx(indices)] = positions

    def set_joint_positions(self, positions: torch.Tensor, indices=None) -> None:
        self.dof_pos[self._idx(indices)] = positions

    def set_joint_velocities(self, velocities: torch.Tensor, indices=None) -> None:
        self.dof_vel[self._idx(indices)] = velocities

    def set_world_poses(self, positions=None, orientations=None, indices=None) -> None:
        if positions is not None:
            self.root_pos[self._idx(indices)] = positions
        if orientations is not None:
            self.root_quat[self._idx(indices)] = orientations

    def set_velocities(self, velocities: torch.Tensor, indices=None) -> None:
        self.root_vel[self._idx(indices)] = velocities
