"""Randomizer -- observation / action noise of the reference (utils/domain_randomization/randomize.py:39-305) on the
B200 kernel ``oige_dr_apply_f32`` (csrc/domain_randomization.cu).

Same class name, constructor, attributes and hook names as the reference, so ``VecEnvRLGames.step`` calls it the same
way (envs/vec_env_rlgames.py:59-60,70-71):

    actions = task._dr_randomizer.apply_actions_randomization(actions=actions, reset_buf=task.reset_buf)
    obs     = task._dr_randomizer.apply_observations_randomization(observations=obs, reset_buf=task.reset_buf)

Each hook is ONE launch working in place on the buffer: the reference's ``reset_buf.nonzero()`` host sync, counter
bookkeeping, RNG launches and gathers/scatters are folded into it.  Noise comes from in-kernel Philox4x32-10 keyed by
(seed, GLOBAL env id, step counter, stream) -- independent of how envs are sharded over GPUs; a caller that wants the
reference's exact torch draws passes them as ``noise_reset`` / ``noise_interval`` ([N, dim], read at the flagged envs).

Only the ``observations`` and ``actions`` keys of ``randomization_params`` are part of this build.  ``simulation``,
``rigid_prim_views`` and ``articulation_views`` (gravity, masses, materials, ... through omni.replicator.isaac's
physics view, randomize.py:307-578) are rejected with an error: they act on PhysX objects that do not exist here.
"""
import ctypes as C
import math

import numpy as np
import torch

from omniisaacgymenvs_b200 import _lib

_DIST = {"gaussian": _lib.DR_GAUSSIAN, "normal": _lib.DR_GAUSSIAN, "uniform": _lib.DR_UNIFORM,
         "loguniform": _lib.DR_LOGUNIFORM, "log_uniform": _lib.DR_LOGUNIFORM}
_OP = {"additive": _lib.DR_ADDITIVE, "scaling": _lib.DR_SCALING}
_UNSUPPORTED = ("simulation", "rigid_prim_views", "articulation_views")


def _noise_spec(block, what: str) -> _lib.OigeNoiseSpec:
    if block is None:
        return _lib.OigeNoiseSpec(0, 0, 0, 0, 0.0, 0.0)
    # the reference only prints for unknown names (randomize.py:265,272) and then fails on an undefined variable
    if block["distribution"] not in _DIST:
        raise ValueError(f"The specified {block['distribution']} distribution is not supported ({what}).")
    if block["operation"] not in _OP:
        raise ValueError(f"The specified {block['operation']} operation type is not supported ({what}).")
    p = block["distribution_parameters"]
    return _lib.OigeNoiseSpec(1, _OP[block["operation"]], _DIST[block["distribution"]],
                              int(block.get("frequency_interval", 1)), float(p[0]), float(p[1]))


class _BufferNoise:
    """Kernel-side state of one randomised buffer (randomize.py:192-193 / :209-210)."""

    def __init__(self, params, num_envs: int, dim: int, device, what: str, stream_id: int):
        self.on_reset = _noise_spec(params.get("on_reset"), f"{what} on_reset")
        self.on_interval = _noise_spec(params.get("on_interval"), f"{what} on_interval")
        self.counter = torch.zeros(num_envs, dtype=torch.int, device=device)
        self.correlated = torch.zeros((num_envs, dim), device=device)
        self.state = _lib.OigeNoiseState(self.counter.data_ptr(), self.correlated.data_ptr())
        self.dim, self.stream_id, self.what = dim, stream_id, what


class Randomizer():
    def __init__(self, sim_config):
        self._cfg = sim_config.task_config
        self._config = sim_config.config

        self.randomize = False
        dr_config = self._cfg.get("domain_randomization", None)
        self.distributions = dict()
        self.active_domain_randomizations = dict()
        self._observations_dr_params = None
        self._actions_dr_params = None
        self._obs_noise = self._act_noise = None
        self._seed = int(self._config.get("seed", 42)) & 0xFFFFFFFFFFFFFFFF
        self._calls = 0           # host mirror of the Philox step counter
        self._counter_dev = None  # device scalar used instead while a CUDA graph replays the step
        self._env_id_offset = int(self._config.get("env_id_offset", 0))

        if dr_config is not None:
            randomize = dr_config.get("randomize", False)
            randomization_params = dr_config.get("randomization_params", None)
            if randomize and randomization_params is not None:
                self.randomize = True
                self.min_frequency = dr_config.get("min_frequency", 1)

    # ---- set-up (randomize.py:126-210) ------------------------------------------------------------------------
    def apply_on_startup_domain_randomization(self, task):
        """randomize.py:59-124 randomises rigid-body scale / mass / density through USD and PhysX views before the
        simulation starts; those objects do not exist in this build."""
        if not self.randomize:
            return self._explain_inactive("On Startup Domain randomization will not be applied.")
        params = self._cfg["domain_randomization"]["randomization_params"]
        for opt in ("rigid_prim_views", "articulation_views"):
            for view_name, attrs in (params.get(opt) or {}).items():
                for attribute, p in (attrs or {}).items():
                    if p is not None and "on_startup" in p:
                        raise RuntimeError(f"on_startup randomisation of {view_name}.{attribute} needs PhysX/USD views and is "
                                           "not part of this build (SURVEY.md section 8(f) N3 covers observation/action noise)")

    def set_up_domain_randomization(self, task):
        if not self.randomize:
            return self._explain_inactive("Domain randomization will not be applied.")
        randomization_params = self._cfg["domain_randomization"]["randomization_params"]
        for opt in randomization_params.keys():
            if opt == "observations":
                self._set_up_observations_randomization(task)
            elif opt == "actions":
                self._set_up_actions_randomization(task)
            elif opt in _UNSUPPORTED and randomization_params[opt] is not None:
                raise RuntimeError(f"domain randomisation of '{opt}' acts on PhysX objects through omni.replicator.isaac and "
                                   "is not part of this build (SURVEY.md section 8(f) N3 covers observation/action noise)")

    def _explain_inactive(self, msg):
        dr_config = self._cfg.get("domain_randomization", None)
        if dr_config is None:
            raise ValueError("No domain randomization parameters are specified in the task yaml config file")
        if dr_config.get("randomize", False) == False or dr_config.get("randomization_params", None) is None:
            print(msg)

    @staticmethod
    def _validate(params, what):
        if params is None:
            raise ValueError(f"{what.capitalize()} randomization parameters are not provided.")
        if "on_reset" in params.keys():
            if not set(('operation', 'distribution', 'distribution_parameters')).issubset(params["on_reset"].keys()):
                raise ValueError(f"Please ensure the following {what} on_reset randomization parameters are provided: "
                                 "operation, distribution, distribution_parameters.")
        if "on_interval" in params.keys():
            if not set(('frequency_interval', 'operation', 'distribution', 'distribution_parameters')).issubset(params["on_interval"].keys()):
                raise ValueError(f"Please ensure the following {what} on_interval randomization parameters are provided: "
                                 "frequency_interval, operation, distribution, distribution_parameters.")

    def _set_up_observations_randomization(self, task):
        task.randomize_observations = True
        self._observations_dr_params = self._cfg["domain_randomization"]["randomization_params"]["observations"]
        self._validate(self._observations_dr_params, "observations")
        for when in ("on_reset", "on_interval"):
            if when in self._observations_dr_params:
                self.active_domain_randomizations[("observations", when)] = np.array(self._observations_dr_params[when]["distribution_parameters"])
        self._obs_noise = _BufferNoise(self._observations_dr_params, task.num_envs, task.num_observations,
                                       self._config["sim_device"], "observations", _lib.DR_STREAM_OBSERVATIONS)
        self._observations_counter_buffer = self._obs_noise.counter
        # the observation pass clamps to +-clipObservations itself; with noise active the clamp has to come AFTER the
        # noise (vec_env_rlgames.py:42 vs :70-71), so it moves into the noise launch (apply_..., clip=task.clip_obs)
        # (task._kernel_clip_obs() keeps returning +inf from now on, so rebuilt parameter structs stay consistent;
        # BOTH the fused-step and the hook-by-hook structs are patched: the reference applies the noise first and
        # clamps once on either path)
        for name in ("_params", "_params_noinc"):
            p = getattr(task, name, None)
            if p is not None and hasattr(p, "clip_obs"):
                p.clip_obs = math.inf
        self._observations_correlated_noise = self._obs_noise.correlated

    def _set_up_actions_randomization(self, task):
        task.randomize_actions = True
        self._actions_dr_params = self._cfg["domain_randomization"]["randomization_params"]["actions"]
        self._validate(self._actions_dr_params, "actions")
        for when in ("on_reset", "on_interval"):
            if when in self._actions_dr_params:
                self.active_domain_randomizations[("actions", when)] = np.array(self._actions_dr_params[when]["distribution_parameters"])
        self._act_noise = _BufferNoise(self._actions_dr_params, task.num_envs, task.num_actions,
                                       self._config["sim_device"], "actions", _lib.DR_STREAM_ACTIONS)
        self._actions_counter_buffer = self._act_noise.counter
        self._actions_correlated_noise = self._act_noise.correlated

    # ---- per-step hooks (randomize.py:212-265) ----------------------------------------------------------------
    def use_device_counter(self, counter_dev: torch.Tensor) -> None:
        """CUDA-graph mode: take the Philox step counter from this device scalar (int64/uint64, advanced by the
        captured step itself) instead of the host-side call count."""
        self._counter_dev = counter_dev

    def _apply(self, bn: _BufferNoise, buffer, reset_buf, clip_in, clip_out, noise_reset, noise_interval, out=None,
               counter=None):
        if bn is None:
            raise RuntimeError("set_up_domain_randomization(task) has not been called for this buffer")
        out = buffer if out is None else out
        for t, nm in ((buffer, "buffer"), (out, "out")):
            if t.dtype != torch.float32 or not t.is_contiguous() or t.shape != (bn.counter.shape[0], bn.dim):
                raise RuntimeError(f"{bn.what}: {nm} must be a contiguous float32 [{bn.counter.shape[0]}, {bn.dim}] tensor")
        self._calls += 1
        rc = _lib.load().oige_dr_apply_f32(
            C.byref(bn.on_reset), C.byref(bn.on_interval), C.byref(bn.state),
            _lib.dptr(buffer, torch.float32, "buffer"), _lib.dptr(out, torch.float32, "out"),
            _lib.dptr(reset_buf, torch.int64, "reset_buf"), buffer.shape[0], bn.dim, float(clip_in), float(clip_out),
            _lib.dptr(noise_reset, torch.float32, "noise_reset"), _lib.dptr(noise_interval, torch.float32, "noise_interval"),
            self._seed, self._calls if counter is None else int(counter), None if self._counter_dev is None else self._counter_dev.data_ptr(),
            bn.stream_id, self._env_id_offset, _lib.current_stream_ptr(buffer.device))
        _lib.check(rc, "oige_dr_apply_f32")
        return out

    def apply_observations_randomization(self, observations, reset_buf, noise_reset=None, noise_interval=None, clip=math.inf,
                                         counter=None):
        """``clip`` folds the wrapper's observation clamp (vec_env_rlgames.py:42, after the noise) into the launch;
        ``counter`` overrides the Philox step counter (default: the number of hook calls so far)."""
        return self._apply(self._obs_noise, observations, reset_buf, math.inf, clip, noise_reset, noise_interval,
                           counter=counter)

    def apply_actions_randomization(self, actions, reset_buf, noise_reset=None, noise_interval=None, clip=math.inf, out=None,
                                    counter=None):
        """``clip`` / ``out`` let the env wrapper fold its own clamp-and-clone of the policy's actions
        (vec_env_rlgames.py:57) into the same launch."""
        return self._apply(self._act_noise, actions, reset_buf, clip, math.inf, noise_reset, noise_interval, out,
                           counter=counter)
