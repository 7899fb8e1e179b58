"""GPU: the drop-in surface (VecEnvRLGames / RLTask / ArticulationView) end to end."""
import numpy as np
import pytest
import torch

import helpers
from omniisaacgymenvs_b200.utils.task_util import make_env
from oracle import locomotion_oracle as lo

pytestmark = pytest.mark.gpu


def _snapshot(task):
    v = task.get_robot()
    return {k: getattr(v, k).clone() for k in ("root_pos", "root_quat", "root_vel", "dof_pos", "dof_vel", "sensor_forces")}


@pytest.mark.parametrize("robot,obs,act", [("Humanoid", 87, 21), ("Ant", 60, 8)])
def test_env_surface_and_shapes(cuda_lib, robot, obs, act):
    env, task = make_env(robot, num_envs=256)
    assert env.num_envs == 256 and task.num_observations == obs and task.num_actions == act
    assert env.observation_space.shape == (obs,) and env.action_space.shape == (act,)
    assert env.num_states == 0 and env.get_number_of_agents() == 1
    d = env.reset()
    assert d["obs"].shape == (256, obs) and d["states"].shape == (256, 0)
    o, r, done, info = env.step(torch.rand(256, act, device=task.device) * 2 - 1)
    assert o["obs"].shape == (256, obs) and r.shape == (256,) and done.shape == (256,) and done.dtype == torch.int64
    assert isinstance(info, dict) and torch.isfinite(o["obs"]).all() and torch.isfinite(r).all()
    v = task.get_robot()
    p, q = v.get_world_poses(clone=False)
    assert p.shape == (256, 3) and q.shape == (256, 4) and v.get_velocities(clone=False).shape == (256, 6)
    assert v.get_joint_positions(clone=False).shape == (256, act)
    assert v._physics_view.get_force_sensor_forces().shape == (256, task._model.num_sensors, 6)
    assert v.get_dof_limits().shape == (256, act, 2) and v.num_dof == act and v.count == 256
    assert p.data_ptr() == v.root_pos.data_ptr(), "clone=False must hand out the live buffer"


@pytest.fixture
def k2_mode(monkeypatch):
    """Pins the articulation kernel (the library reads OIGE_* once: set, then reload)."""
    from omniisaacgymenvs_b200 import _lib

    def pin(mode):
        if mode == "auto":
            monkeypatch.delenv("OIGE_K2_MODE", raising=False)
            monkeypatch.delenv("OIGE_K2_EPC", raising=False)
        else:
            monkeypatch.setenv("OIGE_K2_MODE", "role")
            monkeypatch.setenv("OIGE_K2_EPC", "128" if mode == "role128" else "32")
        _lib.reload_env()

    yield pin
    monkeypatch.undo()
    _lib.reload_env()


@pytest.mark.parametrize("mode,n", [("auto", 512), ("role", 500), ("role128", 1000), ("role128", 1024)])
@pytest.mark.parametrize("robot", ["Humanoid", "Ant"])
def test_fused_step_equals_hook_by_hook_path(cuda_lib, robot, mode, n, k2_mode):
    """oige_env_step_f32 (one call; ONE launch of the whole-step kernel K2-F in the role modes, ragged and full
    tiles) vs pre_physics_step -> world.step x2 -> post_physics_step (K3, the articulation kernel alone, K1)."""
    k2_mode(mode)
    env_a, task_a = make_env(robot, num_envs=n, seed=3)
    env_b, task_b = make_env(robot, num_envs=n, seed=3)
    task_b.fused_step_enabled = lambda: False  # force the generic path
    g = torch.Generator(device="cpu").manual_seed(0)
    env_a.reset(); env_b.reset()
    for step in range(40):
        a = (torch.rand(n, task_a.num_actions, generator=g) * 2.4 - 1.2).to(task_a.device)  # exercises the clamp
        oa, ra, da, _ = env_a.step(a)
        ob, rb, db, _ = env_b.step(a)
        assert torch.equal(oa["obs"], ob["obs"]), f"obs differ at step {step}"
        assert torch.equal(ra, rb) and torch.equal(da, db)
    sa, sb = _snapshot(task_a), _snapshot(task_b)
    for k in sa:
        assert torch.equal(sa[k], sb[k]), k
    assert torch.equal(task_a.progress_buf, task_b.progress_buf)


def test_individual_hooks_equal_fused_post_physics_step(cuda_lib):
    n = 300
    env_a, task_a = make_env("Humanoid", num_envs=n, seed=1)
    env_b, task_b = make_env("Humanoid", num_envs=n, seed=1)
    a = torch.rand(n, 21, device=task_a.device) * 2 - 1
    for env, task in ((env_a, task_a), (env_b, task_b)):
        task.pre_physics_step(a.clone())
        env._world.step(); env._world.step()
    task_a.post_physics_step()
    # reference order, hook by hook (tasks/base/rl_task.py:242-249)
    task_b.progress_buf[:] += 1
    task_b.get_observations(); task_b.calculate_metrics(); task_b.is_done()
    for k in ("obs_buf", "rew_buf", "reset_buf", "progress_buf", "potentials", "prev_potentials", "up_vec", "heading_vec"):
        assert torch.equal(getattr(task_a, k), getattr(task_b, k)), k


@pytest.mark.parametrize("robot", ["Humanoid", "Ant"])
def test_post_physics_step_matches_oracle_on_live_state(cuda_lib, robot):
    """After real physics steps, K1's outputs equal the oracle evaluated on the same state tensors."""
    n = 1024
    env, task = make_env(robot, num_envs=n, seed=5)
    env.reset()
    spec = lo.SPECS[robot]
    for _ in range(10):
        a = torch.rand(n, task.num_actions, device=task.device) * 2 - 1
        pot_before = None
        env.step(a)
    v = task.get_robot()
    # state going INTO the next post_physics_step: run physics by hand, then compare K1 vs oracle
    a = torch.rand(n, task.num_actions, device=task.device) * 2 - 1
    task.pre_physics_step(a)
    env._world.step(); env._world.step()
    s = {"pos": v.root_pos, "quat": v.root_quat, "vel": v.root_vel, "dof_pos": v.dof_pos, "dof_vel": v.dof_vel,
         "sensors": v.sensor_forces, "actions": task.actions, "potentials": task.potentials,
         "progress_buf": task.progress_buf, "reset_buf": task.reset_buf, "targets": task.targets,
         "inv_start_rot": task.inv_start_rot, "basis_vec0": task.basis_vec0, "basis_vec1": task.basis_vec1,
         "dof_lo": task.dof_limits_lower, "dof_hi": task.dof_limits_upper}
    s = {k: t.detach().cpu().clone() for k, t in s.items()}
    ref = lo.post_physics_step(spec, s)
    task.post_physics_step()
    got = {"obs": task.obs_buf, "rew": task.rew_buf, "reset_buf": task.reset_buf, "progress_buf": task.progress_buf,
           "potentials": task.potentials, "prev_potentials": task.prev_potentials, "up_vec": task.up_vec,
           "heading_vec": task.heading_vec}
    got = {k: t.detach().cpu() for k, t in got.items()}
    helpers.assert_k1_matches(got, ref, robot)


def test_sharded_envs_equal_rows_of_the_unsharded_run(cuda_lib):
    """env i of a G-way job must not depend on G.  Shard 1 of 2 (run alone on this
    GPU) == rows [N/2, N) of the 1-way run, bit for bit, resets included."""
    n = 512
    env_full, task_full = make_env("Humanoid", num_envs=n, seed=11)
    env_half, task_half = make_env("Humanoid", num_envs=n // 2, seed=11, env_id_offset=n // 2, global_num_envs=n)
    assert torch.equal(task_full.initial_root_pos[n // 2:], task_half.initial_root_pos)
    env_full.reset(); env_half.reset()
    g = torch.Generator().manual_seed(2)
    for step in range(60):
        a = (torch.rand(n, 21, generator=g) * 2 - 1).to(task_full.device)
        of, rf, df, _ = env_full.step(a)
        oh, rh, dh_, _ = env_half.step(a[n // 2:].contiguous())
        assert torch.equal(of["obs"][n // 2:], oh["obs"]), f"step {step}"
        assert torch.equal(rf[n // 2:], rh) and torch.equal(df[n // 2:], dh_)
    assert int(task_full.read_episode_stats(reset=False)[3]) == n * 61


def test_resets_happen_and_episode_stats_accumulate(cuda_lib):
    n = 2048
    env, task = make_env("Humanoid", num_envs=n, seed=0, overrides={"task.env.episodeLength": 20})
    env.reset()
    task.read_episode_stats()
    total_done = 0
    # episodic return, accumulated here the way the reference's observer would see it per env (rlgames_utils.py:51-62)
    running = task.episode_return.double().clone()  # the reset step's reward already belongs to the first episode
    finished = 0.0
    for _ in range(50):
        _, rew, done, _ = env.step(torch.rand(n, 21, device=task.device) * 2 - 1)
        total_done += int(done.sum())
        running += rew.double()
        finished += float(running[done != 0].sum())
        running[done != 0] = 0.0
    st = task.read_episode_stats()
    # slot 0 is the sum of the RETURNS OF FINISHED EPISODES (per-env accumulator in the kernel), not of every reward
    assert abs(float(st[0]) - finished) <= 1e-4 * abs(finished) + 1e-2
    assert torch.allclose(task.episode_return.double(), running, rtol=1e-5, atol=1e-3)
    assert total_done > n  # every env times out at least twice in 50 steps of 20-step episodes
    assert int(st[1]) == total_done and int(st[3]) == 50 * n
    assert 1.0 <= float(st[2]) / float(st[1]) <= 20.0
    assert int(task.progress_buf.max()) <= 20


def test_humanoid_random_policy_falls_and_ant_stays_up(cuda_lib):
    """Plausibility of the stand-in physics: a random-policy humanoid falls below 0.8 m within a few
    hundred steps in most envs; a zero-action ant settles above its 0.31 m termination height."""
    env, task = make_env("Humanoid", num_envs=1024, seed=0)
    env.reset()
    fell = torch.zeros(1024, dtype=torch.bool, device=task.device)
    for _ in range(300):
        _, _, done, _ = env.step(torch.rand(1024, 21, device=task.device) * 2 - 1)
        fell |= (done != 0) & (task.progress_buf < 999)
    assert float(fell.float().mean()) > 0.5
    env, task = make_env("Ant", num_envs=256, seed=0)
    env.reset()
    for _ in range(200):
        _, _, done, _ = env.step(torch.zeros(256, 8, device=task.device))
    z = task.get_robot().root_pos[:, 2]
    assert float(z.min()) > 0.31 and int(task.reset_buf.sum()) == 0


@pytest.mark.parametrize("n", [512, 20000])
def test_step_host_round_trip(cuda_lib, n):
    """n=20000 exercises the chunked multi-stream pipeline (ragged last chunk) of oige_env_step_host_f32."""
    env_a, task_a = make_env("Humanoid", num_envs=n, seed=4)
    env_b, task_b = make_env("Humanoid", num_envs=n, seed=4)
    env_a.reset(); env_b.reset()
    a = torch.rand(n, 21) * 2 - 1
    ah = a.clone().pin_memory()
    obs_h = torch.empty(n, 87).pin_memory()
    rew_h = torch.empty(n).pin_memory()
    rst_h = torch.empty(n, dtype=torch.int64).pin_memory()
    for _ in range(3):
        env_a.step_host(ah, obs_h, rew_h, rst_h)
        torch.cuda.synchronize()
        o, r, d, _ = env_b.step(a.to(task_b.device))
        assert torch.equal(obs_h, o["obs"].cpu()) and torch.equal(rew_h, r.cpu()) and torch.equal(rst_h, d.cpu())
    va, vb = task_a.get_robot(), task_b.get_robot()
    assert torch.equal(va.dof_pos, vb.dof_pos) and torch.equal(va.root_pos, vb.root_pos)


@pytest.mark.parametrize("robot,n", [("Humanoid", 20000), ("Ant", 4096)])
def test_step_host_cuda_graph_equals_device_step(cuda_lib, robot, n):
    """enable_cuda_graph_host(): H2D, the chunk streams' kernels and D2H captured once per host-buffer set; replays
    must be bit-identical to the device-buffer step, incl. reset noise over several reset waves, and may be mixed
    with un-captured step_host calls."""
    ov = {"task.env.episodeLength": 6}
    env_a, task_a = make_env(robot, num_envs=n, seed=4, overrides=ov)
    env_b, task_b = make_env(robot, num_envs=n, seed=4, overrides=ov)
    env_a.reset(); env_b.reset()
    A = task_a.num_actions
    ring = [(torch.rand(n, A) * 2 - 1).pin_memory() for _ in range(3)]
    obs_h = torch.empty(n, task_a.num_observations).pin_memory()
    rew_h = torch.empty(n).pin_memory()
    rst_h = torch.empty(n, dtype=torch.int64).pin_memory()
    for ah in ring[:2]:
        env_a.enable_cuda_graph_host(ah, obs_h, rew_h, rst_h)
    resets = 0
    for t in range(20):
        ah = ring[t % 3]  # slot 2 was never captured: eager host step in between the replays
        ah.copy_(torch.rand(n, A) * 2 - 1)
        env_a.step_host(ah, obs_h, rew_h, rst_h)
        torch.cuda.synchronize()
        o, r, d, _ = env_b.step(ah.to(task_b.device))
        assert torch.equal(obs_h, o["obs"].cpu()) and torch.equal(rew_h, r.cpu()) and torch.equal(rst_h, d.cpu()), t
        resets += int(rst_h.sum())
    assert resets > n and task_a._reset_counter == task_b._reset_counter
    va, vb = task_a.get_robot(), task_b.get_robot()
    assert torch.equal(va.dof_pos, vb.dof_pos) and torch.equal(va.root_pos, vb.root_pos)


def test_cpu_tensors_are_rejected_not_silently_computed(cuda_lib):
    env, task = make_env("Ant", num_envs=64)
    env.reset()
    with pytest.raises(RuntimeError):
        task.reset_idx(torch.arange(4), rand_pos=torch.rand(4, 8), rand_vel=torch.rand(4, 8))


@pytest.mark.parametrize("robot", ["Humanoid", "Ant"])
def test_task_classes_reproduce_the_reference_episode(cuda_lib, robot):
    """OUR LocomotionTask / ArticulationView on the GPU, driven through the episode recorded from the
    VERBATIM reference task classes (tests/golden/episode_*.npz, oracle/ref_task_driver.py): post_reset
    constants, reset_idx with the reference's torch.rand draws (bit-exact), efforts (bit-exact), then the
    fused post_physics_step on the injected state (1e-5, reset flags bit-exact)."""
    import os

    z = np.load(os.path.join(helpers.GOLDEN_DIR, f"episode_{robot.lower()}.npz"))
    ep = {k: torch.from_numpy(z[k]) for k in z.files}
    steps, n = ep["actions"].shape[:2]
    env, task = make_env(robot, num_envs=n, seed=0)
    dev = task.device
    v = task.get_robot()
    # post_reset constants
    assert torch.equal(task.initial_root_pos.cpu(), ep["initial_root_pos"])
    assert torch.equal(task.initial_root_rot.cpu(), ep["initial_root_rot"])
    assert torch.equal(task.joint_gears.cpu(), ep["joint_gears"])
    assert torch.equal(task.motor_effort_ratio.cpu(), ep["motor_effort_ratio"])
    assert torch.equal(task.dof_limits_lower.cpu(), ep["dof_limits_lower"])
    assert torch.equal(task.dof_limits_upper.cpu(), ep["dof_limits_upper"])
    assert task.dt == 1.0 / 60.0
    # post_reset's reset_idx(all) with the reference's draws
    task.reset_buf.fill_(1)
    task.reset_idx(torch.arange(n, device=dev), ep["post_reset_rand_pos"].to(dev).contiguous(),
                   ep["post_reset_rand_vel"].to(dev).contiguous())
    assert torch.equal(v.dof_pos.cpu(), ep["post_reset_dof_pos"]) and torch.equal(v.dof_vel.cpu(), ep["post_reset_dof_vel"])
    assert torch.equal(task.potentials.cpu(), ep["post_reset_potentials"])
    assert int(task.reset_buf.sum()) == 0 and int(task.progress_buf.sum()) == 0
    for t in range(steps):
        ids = task.reset_buf.nonzero(as_tuple=False).squeeze(-1)
        assert torch.equal(ids.cpu(), ep["reset_ids"][t].nonzero(as_tuple=False).squeeze(-1))
        if len(ids):
            task.reset_idx(ids, ep["rand_pos"][t].to(dev)[ids].contiguous(), ep["rand_vel"][t].to(dev)[ids].contiguous())
        for got, key in ((v.dof_pos, "after_reset_dof_pos"), (v.dof_vel, "after_reset_dof_vel"), (v.root_pos, "after_reset_pos"),
                         (v.root_quat, "after_reset_quat"), (v.root_vel, "after_reset_vel"),
                         (task.potentials, "after_reset_potentials")):
            assert torch.equal(got.cpu(), ep[key][t]), f"step {t}: {key}"
        task.actions.copy_(ep["actions"][t].to(dev))
        forces = task.actions * task.joint_gears * task.power_scale  # locomotion.py:108-109
        assert torch.equal(forces.cpu(), ep["efforts"][t])
        v.root_pos.copy_(ep["inj_pos"][t]); v.root_quat.copy_(ep["inj_quat"][t]); v.root_vel.copy_(ep["inj_vel"][t])
        v.dof_pos.copy_(ep["inj_dof_pos"][t]); v.dof_vel.copy_(ep["inj_dof_vel"][t]); v.sensor_forces.copy_(ep["inj_sensors"][t])
        if t == int(ep["progress_override_step"]):
            task.progress_buf[: n // 4] = task._max_episode_length - 2
        task.post_physics_step()
        got = {"obs": task.obs_buf, "rew": task.rew_buf, "reset_buf": task.reset_buf, "progress_buf": task.progress_buf,
               "potentials": task.potentials, "prev_potentials": task.prev_potentials, "up_vec": task.up_vec,
               "heading_vec": task.heading_vec}
        ref = {k: ep[k][t] for k in got}
        helpers.assert_k1_matches({k: x.detach().cpu() for k, x in got.items()}, ref, f"{robot} step {t}")


@pytest.mark.parametrize("robot,n", [("Humanoid", 512), ("Ant", 4099), ("Humanoid", 16384)])
def test_cuda_graph_step_equals_eager_step(cuda_lib, robot, n):
    """enable_cuda_graph(): the whole step captured once and replayed must be bit-identical to the eager
    step, including the reset noise (device-side Philox counter advanced by the step itself)."""
    env_a, task_a = make_env(robot, num_envs=n, seed=9, overrides={"task.env.episodeLength": 12})
    env_b, task_b = make_env(robot, num_envs=n, seed=9, overrides={"task.env.episodeLength": 12})
    env_a.reset(); env_b.reset()
    env_b.enable_cuda_graph()
    g = torch.Generator().manual_seed(0)
    total_resets = 0
    for step in range(30):
        a = (torch.rand(n, task_a.num_actions, generator=g) * 2.4 - 1.2).to(task_a.device)
        oa, ra, da, _ = env_a.step(a)
        ob, rb, db, _ = env_b.step(a)
        assert torch.equal(oa["obs"], ob["obs"]), f"obs differ at step {step}"
        assert torch.equal(ra, rb) and torch.equal(da, db)
        total_resets += int(da.sum())
    assert total_resets > n  # several reset waves went through the graph
    sa, sb = _snapshot(task_a), _snapshot(task_b)
    for k in sa:
        assert torch.equal(sa[k], sb[k]), k
    assert task_a._reset_counter == task_b._reset_counter
    env_b.disable_cuda_graph()
    oa, ra, da, _ = env_a.step(a)
    ob, rb, db, _ = env_b.step(a)
    assert torch.equal(oa["obs"], ob["obs"]) and torch.equal(da, db)


def test_rlgames_adapter_protocol(cuda_lib):
    """RLGPUEnv / RLGPUAlgoObserver (reference utils/rlgames/rlgames_utils.py:38-118) driven the way rl_games'
    a2c play_steps drives them: get_env_info, reset, step, process_infos, after_print_stats."""
    from omniisaacgymenvs_b200.utils.rlgames.rlgames_utils import RLGPUAlgoObserver, RLGPUEnv

    env, task = make_env("Ant", num_envs=128, overrides={"task.env.episodeLength": 8})
    vec = RLGPUEnv("rlgpu", 128, env_creator=lambda **kw: env)
    info = vec.get_env_info()
    assert info["action_space"].shape == (8,) and info["observation_space"].shape == (60,) and "state_space" not in info
    assert vec.get_number_of_agents() == 1
    obs = vec.reset()
    assert obs["obs"].shape == (128, 60)

    class Writer:
        def __init__(self):
            self.rows = []

        def add_scalar(self, k, v, step):
            self.rows.append((k, float(v), step))

    class Algo:
        games_to_track, writer, device = 10, Writer(), task.device

    ob = RLGPUAlgoObserver()
    ob.after_init(Algo)
    # the task's own episode statistics reach the observer through infos["episode"] (every 16 steps, no sync)
    seen = 0
    for _ in range(64):
        o, r, d, infos = vec.step(torch.rand(128, 8, device=task.device) * 2 - 1)
        if "episode" in infos:
            seen += 1
            assert 1.0 <= float(infos["episode"]["mean_length"]) <= 8.0 and float(infos["episode"]["count"]) >= 128
        ob.process_infos(infos, d.nonzero())
    assert seen >= 2 and len(ob.ep_infos) == seen
    ob.after_print_stats(frame=64 * 128, epoch_num=0, total_time=0.1)
    assert {"Episode/mean_return", "Episode/mean_length", "Episode/count"} <= {k for k, _, _ in Algo.writer.rows}
    for _ in range(3):
        o, r, d, infos = vec.step(torch.rand(128, 8, device=task.device) * 2 - 1)
        infos = dict(infos, episode={"len": task.progress_buf.float().mean()}, lr=3e-4)
        ob.process_infos(infos, d.nonzero())
    ob.mean_scores.update(r)
    ob.after_print_stats(frame=384, epoch_num=1, total_time=0.1)
    keys = {k for k, _, _ in Algo.writer.rows}
    assert {"Episode/len", "lr/frame", "scores/mean"} <= keys


def test_random_policy_script_runs(cuda_lib, capsys):
    from omniisaacgymenvs_b200.scripts import random_policy

    assert random_policy.main(["task=Ant", "num_envs=64", "steps=5"]) == 0
    assert "5 steps of 64 envs" in capsys.readouterr().out


def test_whole_step_is_one_launch_from_4096_humanoids(cuda_lib):
    """clamp + masked reset + substeps + observations / reward / done in ONE kernel launch."""
    from omniisaacgymenvs_b200 import _lib

    env, task = make_env("Humanoid", num_envs=4096, seed=2)
    env.reset()
    a = torch.rand(4096, 21, device=task.device) * 2 - 1
    n0, bad0 = _lib.launch_count(), _lib.nonfinite_env_count()
    env.step(a)
    assert _lib.launch_count() - n0 == 1
    assert _lib.nonfinite_env_count() == bad0  # (cumulative per process: other tests blow envs up on purpose)


def test_graft_entry_smoke_runs(cuda_lib, capsys):
    """The driver's smoke() (one whole-step launch + the two oracle checks) must keep passing as the step changes shape."""
    import __graft_entry__ as g

    g.smoke()
    assert "smoke ok" in capsys.readouterr().out


@pytest.mark.parametrize("chunks", ["1", "4"])
def test_step_host_zero_copy_mode_equals_device_step(cuda_lib, monkeypatch, chunks):
    """OIGE_HOST_ZEROCOPY=3 (what bench.py's e2e leg runs): the whole-step kernel reads the actions from, and stores
    observations / rewards / reset flags into, the pinned host buffers itself; results must equal the device-buffer step
    bit for bit, over several reset waves, eager and as a captured graph."""
    monkeypatch.setenv("OIGE_HOST_ZEROCOPY", "3")
    monkeypatch.setenv("OIGE_HOST_CHUNKS", chunks)
    cuda_lib.oige_reload_env()
    try:
        n = 8192 + 96
        ov = {"task.env.episodeLength": 6}
        env_a, task_a = make_env("Humanoid", num_envs=n, seed=4, overrides=ov)
        env_b, task_b = make_env("Humanoid", num_envs=n, seed=4, overrides=ov)
        env_a.reset(); env_b.reset()
        ring = [(torch.rand(n, 21) * 2 - 1).pin_memory() for _ in range(2)]
        obs_h = torch.empty(n, 87).pin_memory(); rew_h = torch.empty(n).pin_memory()
        rst_h = torch.empty(n, dtype=torch.int64).pin_memory()
        env_a.enable_cuda_graph_host(ring[0], obs_h, rew_h, rst_h)  # slot 1 stays eager
        resets = 0
        for t in range(14):
            ah = ring[t % 2]
            ah.copy_(torch.rand(n, 21) * 2 - 1)
            env_a.step_host(ah, obs_h, rew_h, rst_h)
            torch.cuda.synchronize()
            o, r, d, _ = env_b.step(ah.to(task_b.device))
            assert torch.equal(obs_h, o["obs"].cpu()) and torch.equal(rew_h, r.cpu()) and torch.equal(rst_h, d.cpu()), t
            resets += int(rst_h.sum())
        assert resets > n
        assert torch.equal(task_a.get_robot().dof_pos, task_b.get_robot().dof_pos)
    finally:
        monkeypatch.undo()
        cuda_lib.oige_reload_env()
