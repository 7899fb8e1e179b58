"""HumanoidLocomotionTask (reference: tasks/humanoid.py:47-127): 87 observations, 21 actions,
spawn height 1.34, per-joint gears, ``motor_effort_ratio = gears / max(gears)`` and the
|q|>0.98 joint-limit cost (computed inside K1, OIGE_LIMIT_COST_HUMANOID)."""
import torch

from omniisaacgymenvs_b200 import _lib
from omniisaacgymenvs_b200.articulations import ArticulationView
from omniisaacgymenvs_b200.robots.model import build_model
from omniisaacgymenvs_b200.tasks.base.rl_task import RLTask
from omniisaacgymenvs_b200.tasks.shared.locomotion import LocomotionTask


def _solver_overrides(sim_config, actor: str = None) -> dict:
    """Model overrides from the task YAML: the reference's own keys (gravity, friction, the `physx` block's
    max_depenetration_velocity, the per-actor block `sim.<actor>` with enable_self_collisions /
    max_depenetration_velocity, cfg/task/Humanoid.yaml:34-90) and `sim.b200`, the constants of the stand-in solver."""
    sim = sim_config.task_config.get("sim", {})
    out = {}
    if "gravity" in sim:
        out["gravity"] = tuple(float(g) for g in sim["gravity"])
    mat = sim.get("default_physics_material", {})
    if "dynamic_friction" in mat:
        out["friction_mu"] = float(mat["dynamic_friction"])
    physx = sim.get("physx", {}) or {}
    if "max_depenetration_velocity" in physx:
        out["max_depenetration_velocity"] = float(physx["max_depenetration_velocity"])
    block = (sim.get(actor, {}) or {}) if actor else {}
    if "max_depenetration_velocity" in block and float(block["max_depenetration_velocity"]) > 0:
        out["max_depenetration_velocity"] = float(block["max_depenetration_velocity"])  # per-body value wins
    if "enable_self_collisions" in block:
        out["enable_self_collisions"] = bool(block["enable_self_collisions"])
    out.update({k: float(v) for k, v in sim.get("b200", {}).items()})
    return out


class HumanoidLocomotionTask(LocomotionTask):
    limit_cost_kind = _lib.LIMIT_COST_HUMANOID

    def __init__(self, name, sim_config, env, offset=None) -> None:
        self._sim_config = sim_config
        self._cfg = sim_config.config
        self._task_cfg = sim_config.task_config
        self._num_observations = 87
        self._num_actions = 21
        self._humanoid_positions = torch.tensor([0, 0, 1.34])
        LocomotionTask.__init__(self, name=name, env=env)

    def set_up_scene(self, scene) -> None:
        self.get_humanoid()
        RLTask.set_up_scene(self, scene)
        self._humanoids = ArticulationView(self._model, self._num_envs, self._device, self._env_pos,
                                           name="humanoid_view")
        scene.add(self._humanoids)

    def get_humanoid(self):
        self._model = build_model("Humanoid", spawn_height=float(self._humanoid_positions[2]),
                                  **_solver_overrides(self._sim_config, "Humanoid"))

    def get_robot(self):
        return self._humanoids

    def post_reset(self):
        # tasks/humanoid.py:82-112
        self.joint_gears = torch.tensor(self._model.joint_gears, device=self._device, dtype=torch.float32)
        self.max_motor_effort = torch.max(self.joint_gears)
        self.motor_effort_ratio = self.joint_gears / self.max_motor_effort
        dof_limits = self._humanoids.get_dof_limits()
        self.dof_limits_lower = dof_limits[0, :, 0].to(self._device)
        self.dof_limits_upper = dof_limits[0, :, 1].to(self._device)
        LocomotionTask.post_reset(self)
