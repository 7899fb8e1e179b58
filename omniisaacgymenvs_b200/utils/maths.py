"""Host-side helpers with the names the reference imports from ``omni.isaac.core.utils.torch``
(tasks/shared/locomotion.py:34-35, scripts/rlgames_train.py:107).  Only init-time / glue code uses
them; the per-step arithmetic lives in the CUDA kernels."""
import os
import random

import numpy as np
import torch


def set_seed(seed, torch_deterministic: bool = False):
    if seed == -1:
        seed = np.random.randint(0, 10000)
    print("Setting seed: {}".format(seed))
    random.seed(seed)
    np.random.seed(seed)
    torch.manual_seed(seed)
    os.environ["PYTHONHASHSEED"] = str(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed)
    if torch_deterministic:
        torch.backends.cudnn.benchmark = False
        torch.backends.cudnn.deterministic = True
    return seed


def quat_conjugate(q: torch.Tensor) -> torch.Tensor:
    return torch.cat((q[..., 0:1], -q[..., 1:]), dim=-1)


def torch_rand_float(lower, upper, shape, device):
    return (upper - lower) * torch.rand(*shape, device=device) + lower


def tensor_clamp(t, min_t, max_t):
    return torch.max(torch.min(t, max_t), min_t)


def unscale(x, lower, upper):
    return (2.0 * x - upper - lower) / (upper - lower)
