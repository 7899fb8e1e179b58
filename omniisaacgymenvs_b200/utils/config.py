"""YAML configuration without Hydra/OmegaConf.

The reference composes ``cfg/config.yaml`` + ``cfg/task/<Task>.yaml`` with Hydra 1.1 and four
custom OmegaConf resolvers (utils/hydra_cfg/hydra_utils.py:36-41).  Neither library is available
here, so this module loads the same layout with PyYAML and resolves the small interpolation
subset those files use:

    ${a.b}            absolute reference          ${.x} ${..x} ${...x}   relative references
    ${eq:a,b}  ${contains:a,b}  ${if:p,a,b}  ${resolve_default:default,arg}

The result is a plain nested dict shaped like the reference's ``cfg_dict``
(scripts/rlgames_train.py:89-90): top-level run keys plus ``cfg["task"]``.
"""
import copy
import os
from typing import Any, Dict, List, Optional

import yaml

CFG_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "cfg")

_RESOLVERS = {
    "eq": lambda x, y: str(x).lower() == str(y).lower(),
    "contains": lambda x, y: str(x).lower() in str(y).lower(),
    "if": lambda pred, a, b: a if pred else b,
    "resolve_default": lambda default, arg: default if arg == "" else arg,
}


def _split_args(s: str) -> List[str]:
    out, depth, cur = [], 0, ""
    for ch in s:
        if ch == "," and depth == 0:
            out.append(cur)
            cur = ""
            continue
        if ch == "{":
            depth += 1
        elif ch == "}":
            depth -= 1
        cur += ch
    out.append(cur)
    return [a.strip() for a in out]


def _literal(s: str) -> Any:
    s = s.strip()
    if len(s) >= 2 and s[0] == s[-1] and s[0] in "\"'":
        return s[1:-1]
    try:
        return yaml.safe_load(s) if s != "" else ""
    except yaml.YAMLError:
        return s


class _Resolver:
    def __init__(self, root: dict):
        self.root = root

    def lookup(self, path: List[str], ref: str) -> Any:
        """path = keys from the root to the node holding the interpolation."""
        if ref.startswith("."):
            dots = len(ref) - len(ref.lstrip("."))
            base = path[:-1]                      # container of the current key
            base = base[: len(base) - (dots - 1)] if dots > 1 else base
            keys = base + [k for k in ref.lstrip(".").split(".") if k]
        else:
            keys = ref.split(".")
        node: Any = self.root
        for i, k in enumerate(keys):
            if not isinstance(node, dict) or k not in node:
                raise KeyError(f"interpolation ${{{ref}}} at {'.'.join(path)}: key '{k}' not found")
            node = node[k]
            node = self.resolve(node, keys[: i + 1])
        return node

    def expr(self, body: str, path: List[str]) -> Any:
        body = body.strip()
        head = body.split(":", 1)[0]
        if ":" in body and head in _RESOLVERS:
            args = [self.value(a, path) for a in _split_args(body.split(":", 1)[1])]
            return _RESOLVERS[head](*args)
        return self.lookup(path, body)

    def value(self, s: str, path: List[str]) -> Any:
        """Resolves a scalar string that may contain ${...} groups."""
        s = s.strip()
        if "${" not in s:
            return _literal(s)
        # whole-string interpolation keeps the referenced type
        if s.startswith("${") and self._matching(s, 0) == len(s) - 1:
            return self.expr(s[2:-1], path)
        out, i = "", 0
        while i < len(s):
            if s.startswith("${", i):
                j = self._matching(s, i)
                out += str(self.expr(s[i + 2 : j], path))
                i = j + 1
            else:
                out += s[i]
                i += 1
        return out

    @staticmethod
    def _matching(s: str, start: int) -> int:
        depth = 0
        for i in range(start, len(s)):
            if s.startswith("${", i):
                depth += 1
            elif s[i] == "}":
                depth -= 1
                if depth == 0:
                    return i
        raise ValueError(f"unbalanced interpolation in '{s}'")

    def resolve(self, node: Any, path: List[str]) -> Any:
        if isinstance(node, str) and "${" in node:
            return self.value(node, path)
        return node

    def resolve_all(self, node: Any, path: List[str]) -> Any:
        if isinstance(node, dict):
            return {k: self.resolve_all(v, path + [k]) for k, v in node.items()}
        if isinstance(node, list):
            return [self.resolve_all(v, path + [str(i)]) for i, v in enumerate(node)]
        return self.resolve(node, path)


def _set_dotted(d: dict, dotted: str, value: Any) -> None:
    keys = dotted.split(".")
    for k in keys[:-1]:
        d = d.setdefault(k, {})
    d[keys[-1]] = value


def load_config(task: str = "Humanoid", overrides: Optional[Dict[str, Any]] = None, cfg_dir: Optional[str] = None,
                task_yaml: Optional[str] = None) -> dict:
    """Composes config.yaml + task/<task>.yaml, applies ``overrides`` (dotted keys, like the
    reference's CLI ``num_envs=… task.env.episodeLength=…``) and resolves interpolations."""
    cfg_dir = cfg_dir or CFG_DIR
    with open(os.path.join(cfg_dir, "config.yaml")) as f:
        cfg = yaml.safe_load(f)
    cfg.pop("defaults", None)
    cfg.pop("hydra", None)
    with open(task_yaml or os.path.join(cfg_dir, "task", f"{task}.yaml")) as f:
        cfg["task"] = yaml.safe_load(f)
    for k, v in (overrides or {}).items():
        _set_dotted(cfg, k, v)
    cfg = _Resolver(cfg).resolve_all(copy.deepcopy(cfg), [])
    return cfg


class SimConfig:
    """The slice of the reference's SimConfig (utils/config_utils/sim_config.py:37-108) the
    locomotion tasks touch: ``config``, ``task_config``, device sanitising, physics params."""

    def __init__(self, config: Optional[dict] = None):
        self._config = config if config is not None else {}
        self._cfg = self._config.get("task", {})
        self._config.setdefault("test", False)
        self._config.setdefault("headless", True)
        self._config.setdefault("device_id", 0)
        sim = self._cfg.get("sim", {})
        # sim_config.py:81-96: the pipeline decides the device string
        if sim.get("use_gpu_pipeline", True):
            self._config["sim_device"] = f"cuda:{self._config['device_id']}"
        else:
            raise RuntimeError("pipeline=cpu requested: this build has no CPU path (B200 CUDA only)")
        self._sim_params = {"dt": sim.get("dt", 1.0 / 60.0), "gravity": sim.get("gravity", [0.0, 0.0, -9.81]),
                            "add_ground_plane": sim.get("add_ground_plane", True),
                            "default_physics_material": sim.get("default_physics_material", {}),
                            "b200": sim.get("b200", {}), "sim_device": self._config["sim_device"]}

    @property
    def config(self) -> dict:
        return self._config

    @property
    def task_config(self) -> dict:
        return self._cfg

    @property
    def sim_params(self) -> dict:
        return self._sim_params

    def get_physics_params(self) -> dict:
        return dict(self._sim_params)

    def parse_actor_config(self, actor_name: str) -> dict:
        return dict(self._cfg.get("sim", {}).get(actor_name, {}))

    def apply_articulation_settings(self, name, prim, cfg) -> None:  # USD plumbing: nothing to do
        return None
