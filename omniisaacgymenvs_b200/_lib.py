"""ctypes binding of ``liboige_b200.so`` (C ABI declared in ``include/oige_b200.h``).

There is NO CPU fallback: if the shared library is missing or a call fails, a ``RuntimeError`` is
raised.  PyTorch owns every device buffer; this module only passes ``tensor.data_ptr()``.
"""
import ctypes as C
import os
from typing import Optional

import torch

_PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG_DIR, "liboige_b200.so")

MAX_DOF = 32
MAX_LINKS = 32
MAX_SPHERES = 48
MAX_SENSORS = 4
MAX_PAIRS = 96
STATIC_STAGING_FLOATS = 2 * 1024 * 1024  # OIGE_STATIC_STAGING_FLOATS: action floats one oige_env_step_host_static_f32 call may stage
STATS_SLOTS = 2048
ABI_VERSION = 4
LIMIT_COST_HUMANOID = 0
LIMIT_COST_ANT = 1

f32 = C.c_float
i32 = C.c_int32
i64 = C.c_int64
u64 = C.c_uint64
vp = C.c_void_p


class OigeTaskParams(C.Structure):
    _fields_ = [
        ("num_dof", i32), ("num_sensors", i32), ("num_obs", i32), ("limit_cost_kind", i32),
        ("walk_target_uses_z", i32), ("progress_increment", i32), ("closed_form_angles", i32), ("reserved_i", i32 * 1),
        ("dt", f32), ("dof_vel_scale", f32), ("angular_velocity_scale", f32), ("contact_force_scale", f32),
        ("heading_weight", f32), ("up_weight", f32), ("actions_cost_scale", f32), ("energy_cost_scale", f32),
        ("joints_at_limit_cost_scale", f32), ("death_cost", f32), ("termination_height", f32),
        ("alive_reward_scale", f32), ("max_episode_length", f32), ("clip_obs", f32), ("power_scale", f32),
        ("reserved_f", f32),
        ("targets", f32 * 4), ("inv_start_rot", f32 * 4), ("basis_vec0", f32 * 4), ("basis_vec1", f32 * 4),
        ("dof_limits_lower", f32 * MAX_DOF), ("dof_limits_upper", f32 * MAX_DOF),
        ("motor_effort_ratio", f32 * MAX_DOF), ("joint_gears", f32 * MAX_DOF),
    ]


class OigeState(C.Structure):
    _fields_ = [("root_pos", vp), ("root_quat", vp), ("root_vel", vp), ("dof_pos", vp), ("dof_vel", vp),
                ("sensors", vp), ("dof_pos_targets", vp)]


class OigeTaskBuffers(C.Structure):
    _fields_ = [("obs", vp), ("rew", vp), ("reset_buf", vp), ("progress_buf", vp), ("potentials", vp),
                ("prev_potentials", vp), ("up_vec", vp), ("heading_vec", vp), ("actions", vp), ("episode_stats", vp), ("reset_counter", vp),
                ("episode_return", vp)]


class OigeResetInit(C.Structure):
    _fields_ = [("initial_root_pos", vp), ("initial_root_rot", vp), ("initial_dof_pos", vp)]


class OigeModelDesc(C.Structure):
    _fields_ = [
        ("num_links", i32), ("num_spheres", i32), ("num_sensors", i32), ("fixed_base", i32),
        ("parent", i32 * MAX_LINKS), ("dof_index", i32 * MAX_LINKS), ("joint_type", i32 * MAX_LINKS),
        ("joint_pos", (f32 * 3) * MAX_LINKS), ("joint_quat", (f32 * 4) * MAX_LINKS),
        ("joint_axis", (f32 * 3) * MAX_LINKS), ("mass", f32 * MAX_LINKS), ("com", (f32 * 3) * MAX_LINKS),
        ("inertia", (f32 * 6) * MAX_LINKS), ("armature", f32 * MAX_LINKS), ("damping", f32 * MAX_LINKS),
        ("stiffness", f32 * MAX_LINKS), ("spring_ref", f32 * MAX_LINKS), ("limit_lower", f32 * MAX_LINKS),
        ("limit_upper", f32 * MAX_LINKS), ("drive_stiffness", f32 * MAX_LINKS), ("drive_damping", f32 * MAX_LINKS),
        ("drive_max_force", f32 * MAX_LINKS),
        ("sphere_link", i32 * MAX_SPHERES), ("sphere_pos", (f32 * 3) * MAX_SPHERES),
        ("sphere_radius", f32 * MAX_SPHERES), ("sensor_link", i32 * MAX_SENSORS),
        ("gravity", f32 * 3), ("contact_stiffness", f32), ("contact_damping", f32), ("friction_mu", f32),
        ("friction_viscous", f32), ("limit_stiffness", f32), ("limit_damping", f32), ("max_joint_velocity", f32),
        ("angular_damping", f32),
        ("max_angular_velocity", f32), ("max_depenetration_velocity", f32), ("self_contact_stiffness", f32),
        ("self_contact_damping", f32), ("num_collision_pairs", i32), ("pair_link_a", i32 * MAX_PAIRS),
        ("pair_link_b", i32 * MAX_PAIRS),
        ("capsule_a", (f32 * 3) * MAX_LINKS), ("capsule_b", (f32 * 3) * MAX_LINKS), ("capsule_radius", f32 * MAX_LINKS),
    ]


class OigeCartpoleParams(C.Structure):
    _fields_ = [("cart_dof", i32), ("pole_dof", i32), ("progress_increment", i32), ("reserved_i", i32),
                ("reset_dist", f32), ("max_push_effort", f32), ("max_episode_length", f32), ("clip_obs", f32)]


class OigeCartpoleBuffers(C.Structure):
    _fields_ = [("obs", vp), ("rew", vp), ("reset_buf", vp), ("progress_buf", vp), ("efforts", vp),
                ("episode_stats", vp), ("reset_counter", vp)]


ANYMAL_DOF = 12
ANYMAL_OBS = 48


class OigeAnymalParams(C.Structure):
    _fields_ = [("progress_increment", i32), ("reserved_i", i32 * 3),
                ("lin_vel_scale", f32), ("ang_vel_scale", f32), ("dof_pos_scale", f32), ("dof_vel_scale", f32),
                ("action_scale", f32), ("dt", f32),
                ("rew_lin_vel_xy", f32), ("rew_ang_vel_z", f32), ("rew_lin_vel_z", f32), ("rew_joint_acc", f32),
                ("rew_action_rate", f32), ("rew_cosmetic", f32),
                ("command_x", f32 * 2), ("command_y", f32 * 2), ("command_yaw", f32 * 2),
                ("fallen_threshold", f32), ("max_episode_length", f32), ("clip_obs", f32), ("reserved_f", f32),
                ("default_dof_pos", f32 * ANYMAL_DOF), ("dof_limits_lower", f32 * ANYMAL_DOF),
                ("dof_limits_upper", f32 * ANYMAL_DOF)]


class OigeAnymalBuffers(C.Structure):
    _fields_ = [("obs", vp), ("rew", vp), ("reset_buf", vp), ("progress_buf", vp), ("actions", vp),
                ("last_actions", vp), ("last_dof_vel", vp), ("commands", vp), ("current_targets", vp),
                ("episode_stats", vp), ("reset_counter", vp)]


DR_GAUSSIAN, DR_UNIFORM, DR_LOGUNIFORM = 0, 1, 2
DR_ADDITIVE, DR_SCALING = 0, 1
DR_STREAM_ACTIONS, DR_STREAM_OBSERVATIONS = 0, 2


class OigeNoiseSpec(C.Structure):
    _fields_ = [("enabled", i32), ("operation", i32), ("distribution", i32), ("frequency_interval", i32),
                ("p0", f32), ("p1", f32)]


class OigeNoiseState(C.Structure):
    _fields_ = [("counter", vp), ("correlated", vp)]


# every symbol include/oige_b200.h declares (tests/test_capi_symbols.py checks the .so exports them)
EXPORTED_SYMBOLS = [
    "oige_abi_version", "oige_last_error", "oige_kernel_launch_count", "oige_host_trace_dump",
    "oige_reload_env", "oige_nonfinite_env_count",
    "oige_obs_reward_reset_f32", "oige_obs_reward_reset_tile_f32",
    "oige_reset_masked_f32", "oige_reset_idx_f32",
    "oige_model_create", "oige_model_destroy", "oige_articulation_step_f32",
    "oige_env_step_f32", "oige_env_step_host_f32", "oige_env_step_host_static_f32",
    "oige_cartpole_obs_reward_reset_f32", "oige_cartpole_reset_masked_f32", "oige_cartpole_reset_idx_f32",
    "oige_cartpole_env_step_f32",
    "oige_anymal_obs_reward_reset_f32", "oige_anymal_reset_masked_f32", "oige_anymal_reset_idx_f32",
    "oige_anymal_env_step_f32",
    "oige_dr_apply_f32",
]

_lib: Optional[C.CDLL] = None


def load() -> C.CDLL:
    """Loads the CUDA library; raises RuntimeError (never falls back) if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: the CUDA extension is required (no CPU fallback). "
            "Build it with `python -m omniisaacgymenvs_b200.build`."
        )
    lib = C.CDLL(LIB_PATH)
    P = C.POINTER
    lib.oige_abi_version.restype = C.c_int
    lib.oige_last_error.restype = C.c_char_p
    lib.oige_kernel_launch_count.restype = u64
    lib.oige_nonfinite_env_count.restype = u64
    lib.oige_reload_env.restype = None
    lib.oige_obs_reward_reset_f32.argtypes = [P(OigeTaskParams), P(OigeState), P(OigeTaskBuffers), i64, vp]
    lib.oige_obs_reward_reset_tile_f32.argtypes = [P(OigeTaskParams), P(OigeState), P(OigeTaskBuffers), i64, i32, vp]
    lib.oige_reset_masked_f32.argtypes = [P(OigeTaskParams), P(OigeState), P(OigeTaskBuffers), P(OigeResetInit),
                                          vp, vp, u64, u64, i64, i64, vp]
    lib.oige_reset_idx_f32.argtypes = [P(OigeTaskParams), P(OigeState), P(OigeTaskBuffers), P(OigeResetInit),
                                       vp, i64, vp, vp, u64, u64, i64, vp]
    lib.oige_model_create.argtypes = [P(OigeModelDesc)]
    lib.oige_model_create.restype = vp
    lib.oige_model_destroy.argtypes = [vp]
    lib.oige_model_destroy.restype = None
    lib.oige_articulation_step_f32.argtypes = [vp, P(OigeTaskParams), P(OigeState), vp, i64, i32, f32, vp]
    lib.oige_env_step_f32.argtypes = [vp, P(OigeTaskParams), P(OigeState), P(OigeTaskBuffers), P(OigeResetInit),
                                      vp, f32, u64, u64, i64, i64, i32, f32, vp]
    lib.oige_env_step_host_f32.argtypes = [vp, P(OigeTaskParams), P(OigeState), P(OigeTaskBuffers),
                                           P(OigeResetInit), vp, vp, vp, vp, vp, f32, u64, u64, i64, i64, i32,
                                           f32, vp]
    lib.oige_env_step_host_static_f32.argtypes = [vp, P(OigeTaskParams), P(OigeState), P(OigeTaskBuffers),
                                                  P(OigeResetInit), vp, vp, vp, vp, f32, u64, u64, i64, i64, i32,
                                                  f32, vp]
    lib.oige_env_step_host_static_f32.restype = C.c_int
    CP, CB = P(OigeCartpoleParams), P(OigeCartpoleBuffers)
    lib.oige_cartpole_obs_reward_reset_f32.argtypes = [CP, P(OigeState), CB, i64, vp]
    lib.oige_cartpole_reset_masked_f32.argtypes = [CP, P(OigeState), CB, vp, f32, vp, u64, u64, i64, i64, vp]
    lib.oige_cartpole_reset_idx_f32.argtypes = [CP, P(OigeState), CB, vp, i64, vp, u64, u64, i64, vp]
    lib.oige_cartpole_env_step_f32.argtypes = [vp, CP, P(OigeState), CB, vp, f32, u64, u64, i64, i64, i32, f32, vp]
    AP, AB = P(OigeAnymalParams), P(OigeAnymalBuffers)
    lib.oige_anymal_obs_reward_reset_f32.argtypes = [AP, P(OigeState), AB, i64, vp]
    lib.oige_anymal_reset_masked_f32.argtypes = [AP, P(OigeState), AB, P(OigeResetInit), vp, f32, vp, vp, u64, u64, i64,
                                                 i64, vp]
    lib.oige_anymal_reset_idx_f32.argtypes = [AP, P(OigeState), AB, P(OigeResetInit), vp, i64, vp, vp, u64, u64, i64, vp]
    lib.oige_anymal_env_step_f32.argtypes = [vp, AP, P(OigeState), AB, P(OigeResetInit), vp, f32, u64, u64, i64, i64,
                                             i32, f32, vp]
    lib.oige_dr_apply_f32.argtypes = [P(OigeNoiseSpec), P(OigeNoiseSpec), P(OigeNoiseState), vp, vp, vp, i64, i32, f32,
                                      f32, vp, vp, u64, u64, vp, i32, i64, vp]
    lib.oige_dr_apply_f32.restype = C.c_int
    for name in ("oige_anymal_obs_reward_reset_f32", "oige_anymal_reset_masked_f32", "oige_anymal_reset_idx_f32",
                 "oige_anymal_env_step_f32"):
        getattr(lib, name).restype = C.c_int
    for name in ("oige_obs_reward_reset_f32", "oige_obs_reward_reset_tile_f32", "oige_reset_masked_f32",
                 "oige_reset_idx_f32", "oige_articulation_step_f32", "oige_env_step_f32", "oige_env_step_host_f32",
                 "oige_cartpole_obs_reward_reset_f32", "oige_cartpole_reset_masked_f32",
                 "oige_cartpole_reset_idx_f32", "oige_cartpole_env_step_f32"):
        getattr(lib, name).restype = C.c_int
    if lib.oige_abi_version() != ABI_VERSION:
        raise RuntimeError(f"liboige_b200.so ABI version {lib.oige_abi_version()} != {ABI_VERSION}: rebuild the library")
    _lib = lib
    return lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().oige_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{what} failed (code {rc}): {msg}")


def launch_count() -> int:
    return int(load().oige_kernel_launch_count())


def reload_env() -> None:
    """Re-reads the OIGE_* environment knobs (the library reads them once, at first use)."""
    load().oige_reload_env()


def nonfinite_env_count() -> int:
    """Envs parked by the articulation kernels' non-finite guard on the current device since load (0 when healthy)."""
    return int(load().oige_nonfinite_env_count())


def current_stream_ptr(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def dptr(t: Optional[torch.Tensor], dtype=None, what: str = "tensor") -> Optional[int]:
    """Raw device pointer of a contiguous CUDA tensor (None passes through as NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError(f"{what} must be a CUDA tensor (no CPU fallback); got device {t.device}")
    if not t.is_contiguous():
        raise RuntimeError(f"{what} must be contiguous")
    if dtype is not None and t.dtype != dtype:
        raise RuntimeError(f"{what} must be {dtype}, got {t.dtype}")
    return t.data_ptr()
