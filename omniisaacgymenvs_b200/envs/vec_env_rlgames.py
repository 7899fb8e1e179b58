"""VecEnvRLGames -- the wrapper rl_games drives (reference: envs/vec_env_rlgames.py:36-89).

Same surface (``set_task``, ``step``, ``reset``, ``num_states``, ``state_space``) and the same
per-step order: clamp actions -> pre_physics_step -> control_frequency_inv x world.step ->
post_physics_step -> clamp/copy to rl_device.  When the task is a stock locomotion task the whole
sequence runs as ONE C-ABI call (`oige_env_step_f32`: K3 reset, K2 articulation, K1
obs/reward/reset) with no host sync; any task that overrides the hooks takes the generic path.
"""
import ctypes as C
from datetime import datetime

import torch

from omniisaacgymenvs_b200 import _lib
from omniisaacgymenvs_b200.envs.vec_env_base import VecEnvBase


class VecEnvRLGames(VecEnvBase):
    def _process_data(self):
        # vec_env_rlgames.py:41-46.  The obs clamp already happened inside K1 (clip_obs is a kernel
        # parameter); what is left are the hand-off copies to rl_device.
        dev = self._task.rl_device
        self._obs = self._obs.to(dev).clone()
        self._rew = self._rew.to(dev).clone()
        self._states = torch.clamp(self._states, -self._task.clip_obs, self._task.clip_obs).to(dev).clone()
        self._resets = self._resets.to(dev).clone()
        self._extras = self._extras.copy()

    def set_task(self, task, backend="numpy", sim_params=None, init_sim=True) -> None:
        super().set_task(task, backend, sim_params, init_sim)
        self.num_states = self._task.num_states
        self.state_space = self._task.state_space

    # ---- CUDA-graph mode ------------------------------------------------
    def enable_cuda_graph(self) -> None:
        """Captures the whole step (K3 with the action clamp, K2, K1 and the rl_device hand-off copies) in
        ONE CUDA graph.  After this call ``step`` copies the actions into a static buffer and replays the
        graph: one launch per step instead of three kernels + three copies.  The returned tensors are then
        STATIC buffers that are overwritten by the next step (rl_games copies them into its own rollout
        storage right away); everything else behaves as before.  The Philox reset counter lives on the
        device and is advanced by the step itself, so replays draw fresh reset noise."""
        task = self._task
        if not (getattr(task, "supports_fused_step", False) and task.fused_step_enabled()):
            raise RuntimeError("CUDA-graph mode needs a stock locomotion task (fused step)")
        dev = torch.device(task.device)
        self._g_actions = torch.zeros((self.num_envs, task.num_actions), device=dev)
        dt = self._world.get_physics_dt()
        # replays must continue the host-side counter sequence: the captured call uses *counter_dev
        task._reset_counter_dev.fill_(task._reset_counter + 1)
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        self._graph = torch.cuda.CUDAGraph()
        self.graph_replays = 0
        launches0 = _lib.launch_count()
        saved = task._reset_counter
        if task.randomize_actions or task.randomize_observations:
            task._dr_randomizer.use_device_counter(task._reset_counter_dev)  # replays draw fresh noise
            self._dr_actions_buffer()
        with torch.cuda.graph(self._graph, stream=side):
            self._fused_step_with_noise(self._g_actions, dt, graph_mode=True)
            self._obs, self._rew, self._resets = task.obs_buf, task.rew_buf, task.reset_buf
            self._states = task.get_states()
            self._extras = task.extras
            if torch.device(task.rl_device) == dev:
                # SURVEY.md 8(f) N1, "rl_device hand-off fused": the returned tensors are static buffers in this mode
                # anyway, and obs_buf / rew_buf are written whole (already clamped) by the step and never read by the
                # next one, so they ARE the hand-off buffers -- no clone kernels.  reset_buf is read by the next
                # step's reset pass, so the caller still gets its own copy of it.
                self._resets = self._resets.clone()
                self._states = torch.clamp(self._states, -task.clip_obs, task.clip_obs).clone()
            else:
                self._process_data()
        task._reset_counter = saved  # capture does not execute the step
        self.graph_kernels = _lib.launch_count() - launches0  # kernels of liboige_b200.so inside one replay
        self._g_out = ({"obs": self._obs, "states": self._states}, self._rew, self._resets)
        torch.cuda.current_stream(dev).wait_stream(side)

    def disable_cuda_graph(self) -> None:
        self._graph = None
        if self._task.randomize_actions or self._task.randomize_observations:
            self._task._dr_randomizer.use_device_counter(None)

    # ---- observation / action noise around the fused step (vec_env_rlgames.py:59-60,70-71) -----------
    def _dr_actions_buffer(self):
        if getattr(self, "_dr_actions", None) is None:
            self._dr_actions = torch.empty((self.num_envs, self._task.num_actions), device=self._task.device)
        return self._dr_actions

    def _fused_step_with_noise(self, actions, dt, graph_mode=False) -> None:
        """clamp -> action noise -> [K3 K2 K1] -> observation noise -> clamp; each noise hook is one launch that also
        performs the wrapper's clamp, so K3 / K1 run with their own clamps disabled.  The Philox step counter is
        the task's reset counter (host value in eager mode, device scalar in graph mode: identical sequences)."""
        task = self._task
        dr = task._dr_randomizer
        if task.randomize_actions:
            a = actions if actions.dtype == torch.float32 and actions.is_contiguous() else actions.float().contiguous()
            noisy = dr.apply_actions_randomization(actions=a, reset_buf=task.reset_buf, clip=task.clip_actions,
                                                   out=self._dr_actions_buffer(), counter=task._reset_counter + 1)
            task.fused_step(noisy, dt, graph_mode=graph_mode, clip_actions=float("inf"))
        else:
            task.fused_step(actions, dt, graph_mode=graph_mode)
        if task.randomize_observations:
            dr.apply_observations_randomization(observations=task.obs_buf, reset_buf=task.reset_buf,
                                                clip=task.clip_obs, counter=task._reset_counter + 1)

    def step(self, actions):
        task = self._task
        if getattr(self, "_graph", None) is not None:
            self._g_actions.copy_(actions, non_blocking=True)
            self._graph.replay()
            self.graph_replays += 1
            task._reset_counter += 1  # host mirror of the device-side counter
            self.sim_frame_count += task.control_frequency_inv
            self._world.current_time_step_index += task.control_frequency_inv
            if hasattr(task, "poll_episode_extras"):
                task.poll_episode_extras()
            return self._g_out[0], self._g_out[1], self._g_out[2], task.extras.copy()
        if getattr(task, "supports_fused_step", False) and task.fused_step_enabled():
            actions = actions.to(task.device)
            self._fused_step_with_noise(actions, self._world.get_physics_dt())
            self.sim_frame_count += task.control_frequency_inv
            self._world.current_time_step_index += task.control_frequency_inv
            if hasattr(task, "poll_episode_extras"):
                task.poll_episode_extras()
            self._obs, self._rew, self._resets, self._extras = task.obs_buf, task.rew_buf, task.reset_buf, task.extras
        else:
            actions = torch.clamp(actions, -task.clip_actions, task.clip_actions).to(task.device).clone()
            if task.randomize_actions:
                actions = task._dr_randomizer.apply_actions_randomization(actions=actions, reset_buf=task.reset_buf)
            task.pre_physics_step(actions)
            for _ in range(task.control_frequency_inv):
                self._world.step(render=self._render)
                self.sim_frame_count += 1
            self._obs, self._rew, self._resets, self._extras = task.post_physics_step()
            if task.randomize_observations:
                self._obs = task._dr_randomizer.apply_observations_randomization(
                    observations=self._obs, reset_buf=task.reset_buf, clip=task.clip_obs)
        self._states = task.get_states()
        self._process_data()
        obs_dict = {"obs": self._obs, "states": self._states}
        return obs_dict, self._rew, self._resets, self._extras

    def reset(self):
        """Flags every env for reset and steps once with zero actions (vec_env_rlgames.py:80-89)."""
        now = datetime.now().strftime("%Y-%m-%d %H:%M:%S")
        print(f"[{now}] Running RL reset")
        self._task.reset()
        actions = torch.zeros((self.num_envs, self._task.num_actions), device=self._task.device)
        obs_dict, _, _, _ = self.step(actions)
        return obs_dict

    def enable_cuda_graph_host(self, actions_host, obs_host, rew_host, reset_host) -> None:
        """Captures ``step_host`` for THIS set of pinned host buffers (the copy engines' H2D / D2H transfers, the four
        chunk streams of `oige_env_step_host_f32` and their kernels) in one CUDA graph; later ``step_host`` calls
        with the same four buffers replay it -- one launch instead of ~40 stream operations per step.  Several
        buffer sets may be registered (e.g. a ring of action buffers)."""
        task = self._task
        for t in (actions_host, obs_host, rew_host, reset_host):
            if t.is_cuda or not t.is_pinned():
                raise RuntimeError("enable_cuda_graph_host needs pinned host tensors")
        if getattr(self, "_graph", None) is not None:
            raise RuntimeError("device-buffer CUDA-graph mode is active; disable_cuda_graph() first")
        dev = torch.device(task.device)
        if not hasattr(self, "_host_graphs"):
            self._host_graphs = {}
            # creates the staging buffer and the library's streams / events without stepping the envs
            task.fused_step_host(actions_host, obs_host, rew_host, reset_host, self._world.get_physics_dt(), run=False)
            torch.cuda.synchronize(dev)
        task._reset_counter_dev.fill_(task._reset_counter + 1)
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        g = torch.cuda.CUDAGraph()
        saved = task._reset_counter
        with torch.cuda.graph(g, stream=side):
            task.fused_step_host(actions_host, obs_host, rew_host, reset_host, self._world.get_physics_dt(), graph_mode=True)
        task._reset_counter = saved  # capture does not execute the step
        torch.cuda.current_stream(dev).wait_stream(side)
        self._host_graphs[(actions_host.data_ptr(), obs_host.data_ptr(), rew_host.data_ptr(), reset_host.data_ptr())] = g

    def disable_cuda_graph_host(self) -> None:
        self._host_graphs = {}

    # ---- host-buffer entry point (what a CPU-side trainer would call) -------------------------
    def step_host(self, actions_host: torch.Tensor, obs_host: torch.Tensor, rew_host: torch.Tensor,
                  reset_host: torch.Tensor) -> None:
        """Whole step with HOST (ideally pinned) buffers through `oige_env_step_host_f32`:
        H2D actions, K3+K2+K1, D2H obs/rew/resets, all asynchronous on the current stream."""
        if self._task.randomize_actions or self._task.randomize_observations:
            raise RuntimeError("step_host does not apply observation / action noise; use step()")
        g = getattr(self, "_host_graphs", {}).get(
            (actions_host.data_ptr(), obs_host.data_ptr(), rew_host.data_ptr(), reset_host.data_ptr()))
        if g is not None:
            g.replay()
            self._task._reset_counter += 1  # host mirror of the device-side counter
            self.sim_frame_count += self._task.control_frequency_inv
            return
        self._task.fused_step_host(actions_host, obs_host, rew_host, reset_host, self._world.get_physics_dt())
        if getattr(self, "_host_graphs", None):  # keep the device-side counter of the captured graphs in step
            self._task._reset_counter_dev.fill_(self._task._reset_counter + 1)
        self.sim_frame_count += self._task.control_frequency_inv
