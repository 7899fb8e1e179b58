"""GPU: `oige_env_step_host_static_f32`, the host-buffer env step that allocates nothing.

Who needs it: the caller of the host-buffer step is the part of a trainer that lives in HOST memory (a policy loop
with numpy / pinned arrays behind a ctypes or cgo binding).  It has no CUDA runtime of its own to get the [N,D]
device staging buffer `oige_env_step_host_f32` asks for, and a torch process does not want a raw cudaMalloc in its
step loop either (the caching allocator may hold the HBM already; cudaMalloc stalls the device).  The variant keeps
the staging rows in a fixed 8 MiB array inside the library, so here the test hands it host buffers only.

The file is named so that pytest collects it LAST: `-x` stops at the first failure, and a failure here must not hide
the parity tests.  It has to pass alone in a fresh process (first call on the device = the address look-up) and after
the rest of the suite (which leaves the OIGE_* knobs as it found them; no other test touches the static array).
"""
import ctypes as C

import pytest
import torch

from omniisaacgymenvs_b200 import _lib
from omniisaacgymenvs_b200.utils.task_util import make_env

pytestmark = pytest.mark.gpu


def _step_host_static(lib, env, task, actions_host, obs_host, rew_host, reset_host, n=None):
    """One `oige_env_step_host_static_f32` call through the C ABI, with the task's own structs -- what
    `LocomotionTask.fused_step_host` does for `oige_env_step_host_f32`, minus the staging tensor."""
    task._check_bound()
    n = task._num_envs if n is None else n
    task._reset_counter += 1
    return lib.oige_env_step_host_static_f32(
        task._robots.model_handle, C.byref(task._params), C.byref(task._robots.state_struct), C.byref(task._buffers),
        C.byref(task._reset_init), actions_host.data_ptr(), obs_host.data_ptr(), rew_host.data_ptr(),
        reset_host.data_ptr(), float(task.clip_actions), task._reset_seed, task._reset_counter, task.env_id_offset, n,
        int(task.control_frequency_inv), float(env._world.get_physics_dt()), task._stream())


def _host_buffers(n, obs, pinned=True):
    bufs = [torch.full((n, obs), -123.0), torch.full((n,), -123.0), torch.full((n,), -7, dtype=torch.int64)]
    return [b.pin_memory() for b in bufs] if pinned else bufs


# Humanoid 4096: the one-launch K2-F step; Ant 300: ragged, three launches, and SMALLER than the call before it, so
# the array still holds the Humanoid rows beyond the Ant's when the Ant runs.
@pytest.mark.parametrize("robot,n,pinned", [("Humanoid", 4096, True), ("Ant", 300, True), ("Ant", 300, False)])
def test_static_staging_step_equals_the_allocating_entry_points(cuda_lib, robot, n, pinned):
    """Three calls in a row (each later one finds its predecessor's rows in the array) must give the bits of
    `oige_env_step_host_f32` (caller-allocated staging) and of the device-buffer step -- the tolerance the other
    host-step tests use is bit equality (test_env_gpu.py::test_step_host_round_trip)."""
    _lib.reload_env()
    ov = {"task.env.episodeLength": 3}  # the second call resets envs: the Philox counter path is compared too
    env_s, task_s = make_env(robot, num_envs=n, seed=11, overrides=ov)
    env_h, task_h = make_env(robot, num_envs=n, seed=11, overrides=ov)
    env_d, task_d = make_env(robot, num_envs=n, seed=11, overrides=ov)
    for e in (env_s, env_h, env_d):
        e.reset()
    A, OBS = task_s.num_actions, task_s.num_observations
    out_s, out_h = _host_buffers(n, OBS, pinned), _host_buffers(n, OBS, pinned)
    gen = torch.Generator().manual_seed(5)
    for call in range(3):
        a = torch.rand(n, A, generator=gen) * 4 - 2  # beyond +-1: the clamp is part of the step
        ah = a.clone().pin_memory() if pinned else a.clone()
        rc = _step_host_static(cuda_lib, env_s, task_s, ah, *out_s)
        assert rc == 0, cuda_lib.oige_last_error().decode()
        env_h.step_host(ah, *out_h)  # oige_env_step_host_f32
        torch.cuda.synchronize()
        o, r, d, _ = env_d.step(a.to(task_d.device))  # oige_env_step_f32
        for got, ref, dev in zip(out_s, out_h, (o["obs"], r, d)):
            assert torch.equal(got, ref), (robot, call)
            assert torch.equal(got, dev.cpu()), (robot, call)
    assert int(out_s[2].sum()) > 0, "episodeLength 3: the third call ends with time-outs"
    vs, vh = task_s.get_robot(), task_h.get_robot()
    for k in ("root_pos", "root_quat", "root_vel", "dof_pos", "dof_vel", "sensor_forces"):
        assert torch.equal(getattr(vs, k), getattr(vh, k)), k
    assert torch.equal(task_s.obs_buf, task_h.obs_buf) and torch.equal(task_s.progress_buf, task_h.progress_buf)


def test_static_staging_step_refuses_what_does_not_fit(cuda_lib):
    """Beyond OIGE_STATIC_STAGING_FLOATS the call is refused with a message naming the limit, before it touches a
    buffer (the small buffers passed here would be overrun otherwise); n_envs = 0 is the documented warm-up call."""
    env, task = make_env("Humanoid", num_envs=64, seed=1)
    env.reset()
    ah = torch.zeros(64, 21).pin_memory()
    out = _host_buffers(64, 87)
    before = [b.clone() for b in out]
    cnt = task._reset_counter
    too_many = _lib.STATIC_STAGING_FLOATS // 21 + 1
    rc = _step_host_static(cuda_lib, env, task, ah, *out, n=too_many)
    task._reset_counter = cnt
    assert rc == -4  # OIGE_E_RANGE
    msg = cuda_lib.oige_last_error().decode()
    assert "OIGE_STATIC_STAGING_FLOATS" in msg and str(too_many) in msg and "oige_env_step_host_f32" in msg
    assert _step_host_static(cuda_lib, env, task, ah, *out, n=0) == 0
    task._reset_counter = cnt
    torch.cuda.synchronize()
    for b, b0 in zip(out, before):
        assert torch.equal(b, b0)
    assert _step_host_static(cuda_lib, env, task, ah, *out) == 0  # and the env still steps afterwards
    torch.cuda.synchronize()
    assert torch.isfinite(out[0]).all() and torch.isfinite(out[1]).all()
