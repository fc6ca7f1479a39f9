"""Dataset preparation on the GPU: return_reward_range / modify_reward / compute_mean_std / normalize_states of
/root/reference/algorithms/offline/iql.py:132-139, 344-401 behind the same names and signatures (numpy in, numpy out,
`dataset["rewards"]` modified in place), plus `*_device` variants that stay on torch CUDA tensors.  The arithmetic is in
libiqlb.so (`csrc/dataset_prep.cu`); there is no CPU path."""
import ctypes as C
from typing import Dict, Optional, Tuple

import numpy as np
import torch

from . import _lib

_LOCOMOTION = ("halfcheetah", "hopper", "walker2d")


def _dev(device) -> torch.device:
    dev = torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError("iqlpref_b200 has no CPU path: dataset preparation runs on a CUDA device")
    return dev


def _terminals_u8(terminals, dev) -> torch.Tensor:
    t = torch.as_tensor(np.asarray(terminals) != 0 if not torch.is_tensor(terminals) else terminals != 0)
    return t.to(dev, torch.uint8).contiguous()


def return_reward_range_device(rewards: torch.Tensor, terminals: torch.Tensor, max_episode_steps: int,
                               want_lengths: bool = True):
    """rewards float32 [n], terminals uint8 [n] on a CUDA device -> (min_max float64[2] tensor, trj_lens float32[n])."""
    n = rewards.numel()
    if n == 0:
        raise ValueError("empty dataset")
    dev = rewards.device
    min_max = torch.empty(2, dtype=torch.float64, device=dev)
    trj = torch.empty(n, dtype=torch.float32, device=dev) if want_lengths else None
    lib = _lib.load()
    with _lib.device_ctx(dev):
        _lib.check(lib.iqlb_reward_range(_lib.ptr(rewards), _lib.ptr(terminals), n, int(max_episode_steps),
                                         _lib.ptr(min_max), _lib.ptr(trj), _lib.stream_ptr(dev)), "iqlb_reward_range")
    return min_max, trj


def return_reward_range(dataset: Dict[str, np.ndarray], max_episode_steps: int, device="cuda"):
    """iql.py:344-360: (min episode return, max episode return, trj_lens)."""
    dev = _dev(device)
    r = torch.as_tensor(np.ascontiguousarray(dataset["rewards"], dtype=np.float32).reshape(-1)).to(dev)
    t = _terminals_u8(np.asarray(dataset["terminals"]).reshape(-1), dev)
    mm, trj = return_reward_range_device(r, t, max_episode_steps)
    mm = mm.cpu().numpy()
    if not np.isfinite(mm).all():  # no episode was completed: the reference's min([]) raises
        raise ValueError("min() arg is an empty sequence")
    return float(mm[0]), float(mm[1]), trj.cpu().numpy().astype(np.float64)


def modify_reward_device(rewards: torch.Tensor, terminals: torch.Tensor, env_name: str, normalize_reward: int,
                         max_episode_steps: int = 1000) -> torch.Tensor:
    """In place on a float32 CUDA tensor (iql.py:363-401); environments the reference leaves alone are returned as is."""
    if any(s in env_name for s in _LOCOMOTION):
        kind = 0
    elif "antmaze" in env_name:
        kind = 1
    else:
        return rewards
    lib = _lib.load()
    with _lib.device_ctx(rewards.device):
        _lib.check(lib.iqlb_modify_reward(_lib.ptr(rewards), _lib.ptr(terminals), rewards.numel(), kind,
                                          int(normalize_reward), int(max_episode_steps),
                                          _lib.stream_ptr(rewards.device)), "iqlb_modify_reward")
    return rewards


def modify_reward(dataset: Dict[str, np.ndarray], env_name: str, normalize_reward: int, max_episode_steps: int = 1000,
                  device="cuda") -> None:
    """iql.py:363-401: rewrites dataset["rewards"] in place (same dtype) through the GPU."""
    dev = _dev(device)
    rewards = dataset["rewards"]
    r = torch.as_tensor(np.ascontiguousarray(rewards, dtype=np.float32).reshape(-1)).to(dev)
    t = _terminals_u8(np.asarray(dataset["terminals"]).reshape(-1), dev)
    modify_reward_device(r, t, env_name, normalize_reward, max_episode_steps)
    rewards[...] = r.cpu().numpy().reshape(rewards.shape)


def compute_mean_std_device(states: torch.Tensor, eps: float) -> Tuple[torch.Tensor, torch.Tensor]:
    """states float32 [n, dim] (row-strided views allowed) -> (mean, std + eps), float32 [dim]."""
    if states.ndim != 2 or states.stride(1) != 1:
        raise ValueError("states must be [n, dim] with contiguous rows")
    n, dim = states.shape
    mean = torch.empty(dim, dtype=torch.float32, device=states.device)
    std = torch.empty(dim, dtype=torch.float32, device=states.device)
    lib = _lib.load()
    with _lib.device_ctx(states.device):
        _lib.check(lib.iqlb_state_mean_std(_lib.ptr(states), n, dim, states.stride(0), float(eps), _lib.ptr(mean),
                                           _lib.ptr(std), _lib.stream_ptr(states.device)), "iqlb_state_mean_std")
    return mean, std


def compute_mean_std(states: np.ndarray, eps: float, device="cuda") -> Tuple[np.ndarray, np.ndarray]:
    """iql.py:132-135."""
    x = torch.as_tensor(np.ascontiguousarray(states, dtype=np.float32)).to(_dev(device))
    mean, std = compute_mean_std_device(x, eps)
    return mean.cpu().numpy(), std.cpu().numpy()


def normalize_states_device(states: torch.Tensor, mean: torch.Tensor, std: torch.Tensor) -> torch.Tensor:
    """In place (states - mean) / std on a float32 CUDA tensor [n, dim] (row-strided views allowed)."""
    n, dim = states.shape
    lib = _lib.load()
    with _lib.device_ctx(states.device):
        _lib.check(lib.iqlb_normalize_states(_lib.ptr(states), n, dim, states.stride(0) if n > 1 else dim,
                                             _lib.ptr(mean.to(states.device, torch.float32).contiguous()),
                                             _lib.ptr(std.to(states.device, torch.float32).contiguous()),
                                             _lib.stream_ptr(states.device)), "iqlb_normalize_states")
    return states


def normalize_states(states: np.ndarray, mean: np.ndarray, std: np.ndarray, device="cuda") -> np.ndarray:
    """iql.py:138-139."""
    dev = _dev(device)
    x = torch.as_tensor(np.ascontiguousarray(states, dtype=np.float32)).to(dev)
    normalize_states_device(x, torch.as_tensor(np.asarray(mean, dtype=np.float32)).to(dev),
                            torch.as_tensor(np.asarray(std, dtype=np.float32)).to(dev))
    return x.cpu().numpy()
