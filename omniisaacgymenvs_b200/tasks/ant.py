"""AntLocomotionTask (reference: tasks/ant.py:46-95): 60 observations, 8 actions, spawn height
0.5, gears all 15, ``motor_effort_ratio`` all ones and the count-of-(q>0.99) limit cost
(int64 in the reference, subtracted unscaled; computed inside K1, OIGE_LIMIT_COST_ANT)."""
import torch

from omniisaacgymenvs_b200 import _lib
from omniisaacgymenvs_b200.articulations import ArticulationView
from omniisaacgymenvs_b200.robots.model import build_model
from omniisaacgymenvs_b200.tasks.base.rl_task import RLTask
from omniisaacgymenvs_b200.tasks.humanoid import _solver_overrides
from omniisaacgymenvs_b200.tasks.shared.locomotion import LocomotionTask


class AntLocomotionTask(LocomotionTask):
    limit_cost_kind = _lib.LIMIT_COST_ANT

    def __init__(self, name, sim_config, env, offset=None) -> None:
        self._sim_config = sim_config
        self._cfg = sim_config.config
        self._task_cfg = sim_config.task_config
        self._num_observations = 60
        self._num_actions = 8
        self._ant_positions = torch.tensor([0, 0, 0.5])
        LocomotionTask.__init__(self, name=name, env=env)

    def set_up_scene(self, scene) -> None:
        self.get_ant()
        RLTask.set_up_scene(self, scene)
        self._ants = ArticulationView(self._model, self._num_envs, self._device, self._env_pos, name="ant_view")
        scene.add(self._ants)

    def get_ant(self):
        self._model = build_model("Ant", spawn_height=float(self._ant_positions[2]),
                                  **_solver_overrides(self._sim_config, "Ant"))

    def get_robot(self):
        return self._ants

    def post_reset(self):
        # tasks/ant.py:79-86
        self.joint_gears = torch.tensor(self._model.joint_gears, dtype=torch.float32, device=self._device)
        dof_limits = self._ants.get_dof_limits()
        self.dof_limits_lower = dof_limits[0, :, 0].to(self._device)
        self.dof_limits_upper = dof_limits[0, :, 1].to(self._device)
        self.motor_effort_ratio = torch.ones_like(self.joint_gears, device=self._device)
        LocomotionTask.post_reset(self)
