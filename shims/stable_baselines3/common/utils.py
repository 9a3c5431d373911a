"""stable_baselines3.common.utils: the four helpers the reference reaches (main.py:12, actor_critic_ts.py:7)."""
import random

import numpy as np
import torch as th


def get_device(device="auto"):
    """SB3: "auto" -> cuda when available, else cpu."""
    if device == "cuda" or device == "auto":
        device = "cuda" if th.cuda.is_available() else "cpu"
    device = th.device(device)
    if device.type == "cuda" and device.index is None and th.cuda.is_available():
        device = th.device("cuda", th.cuda.current_device())
    return device


def get_linear_fn(start, end, end_fraction):
    """Linear interpolation between start and end as progress_remaining goes 1 -> 1 - end_fraction, then `end`."""
    def func(progress_remaining):
        if (1 - progress_remaining) > end_fraction:
            return end
        return start + (1 - progress_remaining) * (end - start) / end_fraction
    return func


def obs_as_tensor(obs, device):
    if isinstance(obs, th.Tensor):
        return obs.to(device)
    return th.as_tensor(np.asarray(obs)).to(device)


def set_random_seed(seed, using_cuda=False):
    random.seed(seed)
    np.random.seed(seed)
    th.manual_seed(seed)


def explained_variance(y_pred, y_true):
    var_y = np.var(y_true)
    return np.nan if var_y == 0 else 1 - np.var(y_true - y_pred) / var_y
