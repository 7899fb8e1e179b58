"""rl_games adapter (reference: utils/rlgames/rlgames_utils.py:38-118).

``RLGPUEnv`` is the ``IVecEnv`` rl_games instantiates around the already-built ``VecEnvRLGames``;
``RLGPUAlgoObserver`` forwards ``infos['episode']`` / scalar extras to the trainer's summary writer.
rl_games is not installed on the build machines, so the two base classes fall back to plain
``object`` stand-ins with the same method names: the adapter can then still be driven (and is tested)
by anything that speaks the ``IVecEnv`` protocol, and binds to the real classes when rl_games exists.
"""
import torch

try:  # the real base classes when rl_games is available
    from rl_games.common import vecenv as _vecenv
    from rl_games.common.algo_observer import AlgoObserver as _AlgoObserver

    _IVecEnv = _vecenv.IVecEnv
    HAVE_RL_GAMES = True
except Exception:  # pragma: no cover - exercised on machines without rl_games
    HAVE_RL_GAMES = False

    class _IVecEnv:  # rl_games.common.vecenv.IVecEnv protocol
        def step(self, actions):
            raise NotImplementedError

        def reset(self):
            raise NotImplementedError

        def has_action_masks(self):
            return False

        def get_number_of_agents(self):
            return 1

        def get_env_info(self):
            pass

    class _AlgoObserver:  # rl_games.common.algo_observer.AlgoObserver protocol
        def before_init(self, base_name, config, experiment_name):
            pass

        def after_init(self, algo):
            pass

        def process_infos(self, infos, done_indices):
            pass

        def after_steps(self):
            pass

        def after_clear_stats(self):
            pass

        def after_print_stats(self, frame, epoch_num, total_time):
            pass


class _Mean:
    """Running mean over the last `maxlen` scores (stand-in for rl_games' torch_ext.AverageMeter)."""

    def __init__(self, maxlen):
        self.maxlen, self.vals = maxlen, []

    def update(self, v):
        self.vals.extend(float(x) for x in v.reshape(-1))
        self.vals = self.vals[-self.maxlen:]

    def clear(self):
        self.vals = []

    @property
    def current_size(self):
        return len(self.vals)

    def get_mean(self):
        return sum(self.vals) / max(len(self.vals), 1)


class RLGPUAlgoObserver(_AlgoObserver):
    """Logs env stats next to the algorithm's (rlgames_utils.py:38-91)."""

    def __init__(self):
        pass

    def after_init(self, algo):
        self.algo = algo
        self.mean_scores = _Mean(getattr(algo, "games_to_track", 100))
        self.ep_infos = []
        self.direct_info = {}
        self.writer = getattr(algo, "writer", None)

    def process_infos(self, infos, done_indices):
        assert isinstance(infos, dict), "RLGPUAlgoObserver expects dict info"
        if "episode" in infos:
            self.ep_infos.append(infos["episode"])
        if len(infos) > 0:  # allow direct logging of scalars from the env
            self.direct_info = {}
            for k, v in infos.items():
                if isinstance(v, (float, int)) or (isinstance(v, torch.Tensor) and len(v.shape) == 0):
                    self.direct_info[k] = v

    def after_clear_stats(self):
        self.mean_scores.clear()

    def after_print_stats(self, frame, epoch_num, total_time):
        if self.writer is None:
            return
        if self.ep_infos:
            for key in self.ep_infos[0]:
                vals = []
                for ep_info in self.ep_infos:
                    v = ep_info[key]
                    v = v if isinstance(v, torch.Tensor) else torch.tensor([v], dtype=torch.float32)
                    vals.append(v.reshape(-1).float().cpu())
                self.writer.add_scalar("Episode/" + key, torch.mean(torch.cat(vals)), epoch_num)
            self.ep_infos.clear()
        for k, v in self.direct_info.items():
            self.writer.add_scalar(f"{k}/frame", v, frame)
            self.writer.add_scalar(f"{k}/iter", v, epoch_num)
            self.writer.add_scalar(f"{k}/time", v, total_time)
        if self.mean_scores.current_size > 0:
            m = self.mean_scores.get_mean()
            self.writer.add_scalar("scores/mean", m, frame)
            self.writer.add_scalar("scores/iter", m, epoch_num)
            self.writer.add_scalar("scores/time", m, total_time)


class RLGPUEnv(_IVecEnv):
    """rl_games vec-env around the env returned by ``env_creator`` (rlgames_utils.py:94-118)."""

    def __init__(self, config_name, num_actors, **kwargs):
        creator = kwargs.pop("env_creator", None)
        if creator is None:
            from rl_games.common import env_configurations  # registered by the trainer script

            creator = env_configurations.configurations[config_name]["env_creator"]
        self.env = creator(**kwargs)

    def step(self, action):
        return self.env.step(action)

    def reset(self):
        return self.env.reset()

    def get_number_of_agents(self):
        return self.env.get_number_of_agents()

    def get_env_info(self):
        info = {"action_space": self.env.action_space, "observation_space": self.env.observation_space}
        if self.env.num_states > 0:
            info["state_space"] = self.env.state_space
            print(info["action_space"], info["observation_space"], info["state_space"])
        else:
            print(info["action_space"], info["observation_space"])
        return info
