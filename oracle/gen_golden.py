"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/*.npz from the VERBATIM reference functions.

Run in the build container (needs /root/reference):

    python -m oracle.gen_golden

For every case the inputs of SURVEY.md section 8(d) are generated, then the reference's own
``get_observations`` / ``get_dof_at_limit_cost`` / ``calculate_metrics`` / ``is_done``
(tasks/shared/locomotion.py:194-321, tasks/humanoid.py:120-127, tasks/ant.py:92-95 -- imported
verbatim through oracle/ref_loader.py, third-party helpers from oracle/isaac_math.py) produce the
outputs.  The fixtures travel to the GPU box where /root/reference does not exist.
"""
import math
import os
import sys

import numpy as np
import torch

from oracle import locomotion_oracle as lo
from oracle import ref_loader

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def run_reference(spec: lo.TaskSpec, s: dict) -> dict:
    """One post_physics_step through the verbatim reference functions (rl_task.py:231-251 order)."""
    loco, hum, ant = ref_loader.load_reference()
    n = s["pos"].shape[0]
    mer = spec.motor_effort_ratio()
    progress = s["progress_buf"] + 1
    obs, pot, prev_pot, up_vec, heading_vec = loco.get_observations(
        s["pos"], s["quat"], s["vel"][:, 0:3], s["vel"][:, 3:6], s["dof_pos"], s["dof_vel"], s["targets"],
        s["potentials"], spec.dt, s["inv_start_rot"], s["basis_vec0"], s["basis_vec1"], s["dof_lo"], s["dof_hi"],
        spec.dof_vel_scale, s["sensors"], n, spec.contact_force_scale, s["actions"], spec.angular_velocity_scale)
    if spec.name == "Ant":
        cost = ant.get_dof_at_limit_cost(obs, spec.num_dof)
    else:
        cost = hum.get_dof_at_limit_cost(obs, mer, spec.joints_at_limit_cost_scale)
    rew = loco.calculate_metrics(
        obs, s["actions"], spec.up_weight, spec.heading_weight, pot, prev_pot, spec.actions_cost_scale,
        spec.energy_cost_scale, spec.termination_height, spec.death_cost, spec.num_dof, cost,
        spec.alive_reward_scale, mer)
    reset = loco.is_done(obs, spec.termination_height, s["reset_buf"], progress, spec.max_episode_length)
    return {"obs": obs, "rew": rew, "reset_buf": reset, "progress_buf": progress, "potentials": pot,
            "prev_potentials": prev_pot, "up_vec": up_vec, "heading_vec": heading_vec}


def edge_state(spec: lo.TaskSpec) -> dict:
    """Hand-built adversarial rows: thresholds, gimbal branch, degenerate target."""
    n = 48
    s = lo.synthetic_state(spec, n, 1234)
    d = spec.num_dof
    lo_, hi_ = s["dof_lo"], s["dof_hi"]
    h = spec.termination_height
    z = s["pos"][:, 2]
    # rows 0-3: z exactly at / one ulp around the termination height
    z[0] = h
    z[1] = float(np.nextafter(np.float32(h), np.float32(10)))
    z[2] = float(np.nextafter(np.float32(h), np.float32(-10)))
    z[3] = h + 0.5
    z[4:] = h + 0.3
    # rows 4-7: progress at the episode-length boundary
    s["progress_buf"][4:8] = torch.tensor([996, 997, 998, 999])
    s["progress_buf"][8:] = 5
    # rows 8-11: pitch at the +-pi/2 gimbal branch (|sinp| >= 1) and near it
    r2 = math.sqrt(0.5)
    s["quat"][8] = torch.tensor([r2, 0.0, r2, 0.0])
    s["quat"][9] = torch.tensor([r2, 0.0, -r2, 0.0])
    s["quat"][10] = torch.tensor([0.70710677, 1e-4, 0.70710677, 0.0])
    s["quat"][11] = torch.tensor([1.0, 0.0, 0.0, 0.0])
    # rows 12-13: torso exactly above the target (to_target norm 0 -> clamp 1e-9 path) and very close
    s["pos"][12, 0:2] = torch.tensor([1000.0, 0.0])
    s["pos"][13, 0:2] = torch.tensor([1000.0 - 1e-4, 1e-5])
    # rows 14-21: dof_pos giving q_scaled exactly at / around the 0.98 and 0.99 thresholds
    for row, target in zip(range(14, 22), [0.98, -0.98, 0.99, -0.99, 0.9800001, 0.9899999, 1.0, -1.05]):
        s["dof_pos"][row] = 0.5 * (target * (hi_ - lo_) + hi_ + lo_)
    # rows 22-25: identity / flipped orientations (up_proj = +-1, heading_proj = +-1)
    s["quat"][22] = torch.tensor([1.0, 0.0, 0.0, 0.0])
    s["quat"][23] = torch.tensor([0.0, 1.0, 0.0, 0.0])
    s["quat"][24] = torch.tensor([0.0, 0.0, 0.0, 1.0])
    s["quat"][25] = torch.tensor([0.0, 0.0, 1.0, 0.0])
    # rows 26-29: small tilts around the up_proj 0.93 and heading_proj 0.8 thresholds
    for row, ang, axis in ((26, math.acos(0.93), 1), (27, math.acos(0.9301), 1), (28, math.acos(0.8), 3),
                           (29, math.acos(0.7999), 3)):
        q = torch.zeros(4)
        q[0] = math.cos(ang / 2)
        q[axis] = math.sin(ang / 2)
        s["quat"][row] = q
        s["pos"][row, 0:2] = 0.0
    # rows 30-31: zero actions / saturated actions; row 32: reset flag already set on input
    s["actions"][30] = 0.0
    s["actions"][31] = torch.where(torch.arange(d) % 2 == 0, torch.ones(d), -torch.ones(d))
    s["reset_buf"][32] = 1
    # rows 33-35: large velocities and sensor readings
    s["vel"][33] = 100.0
    s["dof_vel"][34] = -50.0
    s["sensors"][35] = 1e4
    return s


def save_case(name: str, spec: lo.TaskSpec, s: dict) -> None:
    out = run_reference(spec, {k: v.clone() for k, v in s.items()})
    blob = {f"in_{k}": v.numpy() for k, v in s.items()}
    blob.update({f"out_{k}": v.numpy() for k, v in out.items()})
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    np.savez_compressed(os.path.join(GOLDEN_DIR, name + ".npz"), **blob)


def write_dr() -> None:
    """Observation / action noise: the verbatim Randomizer methods
    (utils/domain_randomization/randomize.py:212-305); `python -m oracle.gen_golden dr` writes this fixture only."""
    from oracle import ref_dr_driver

    np.savez_compressed(os.path.join(GOLDEN_DIR, "episode_dr.npz"), **ref_dr_driver.run_all())


def main() -> None:
    if sys.argv[1:] == ["dr"]:
        return write_dr()
    if not ref_loader.reference_available():
        raise SystemExit("reference tree not found: golden vectors can only be generated in the build container")
    torch.set_num_threads(1)
    for spec in (lo.HUMANOID, lo.ANT):
        for seed in (0, 1, 2, 3):
            for n in (1, 33, 256):
                save_case(f"{spec.name.lower()}_s{seed}_n{n}", spec, lo.synthetic_state(spec, n, seed))
        save_case(f"{spec.name.lower()}_edge", spec, edge_state(spec))
    # method-level episodes through the verbatim task CLASSES (post_reset / pre_physics_step / reset_idx / ...)
    from omniisaacgymenvs_b200.articulations import grid_env_origins
    from omniisaacgymenvs_b200.robots.model import build_model
    from oracle import ref_task_driver

    for robot in ("Humanoid", "Ant"):
        model = build_model(robot)  # the limits / spawn height are INPUT data of the episode
        n = 96
        spawn = torch.from_numpy(grid_env_origins(n, 5.0))
        spawn[:, 2] += model.spawn_height
        ep = ref_task_driver.run_episode(robot, n, steps=5, seed=0, dof_limits=torch.from_numpy(model.dof_limits()),
                                         spawn_pos=spawn)
        np.savez_compressed(os.path.join(GOLDEN_DIR, f"episode_{robot.lower()}.npz"), **ep)
    # Cartpole: the verbatim CartpoleTask methods on a fake view (tasks/cartpole.py:80-162)
    from oracle import ref_cartpole_driver

    np.savez_compressed(os.path.join(GOLDEN_DIR, "episode_cartpole.npz"), **ref_cartpole_driver.run_episode(96, steps=6, seed=0))
    # Anymal: the verbatim AnymalTask / AnymalView methods on a fake view (tasks/anymal.py:137-293)
    from oracle import ref_anymal_driver

    model = build_model("Anymal")
    spawn = torch.from_numpy(grid_env_origins(96, 4.0))
    spawn[:, 2] += model.spawn_height
    ep = ref_anymal_driver.run_episode(96, steps=5, seed=0, spawn_pos=spawn, dof_limits=torch.from_numpy(model.dof_limits()))
    np.savez_compressed(os.path.join(GOLDEN_DIR, "episode_anymal.npz"), **ep)
    write_dr()
    print("wrote", len(os.listdir(GOLDEN_DIR)), "fixtures to", GOLDEN_DIR)


if __name__ == "__main__":
    main()
