"""Reward-model relabelling: mirrors `qlearning_dataset_{mr,bnn,mr_ensemble,pt}` and the loaders of
/root/reference/algorithms/offline/iql.py:691-1390.  Host logic (file discovery, keep mask,
warnings, diagnostics) follows the reference; every forward pass and the mean/CVaR reduction run
in libiqlb.so on the GPU (`iqlb_relabel_mlp`, `iqlb_relabel_pt`).  No CPU path.
"""
import ctypes as C
import glob as _glob
import math
import os
import re
import warnings
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import yaml

from . import _lib
from .reward_models.pref_transformer import PT as RewardPT
from .reward_models.pref_transformer import packed_pt_weights, pt_desc_from_model
from .reward_models.q_mlp import MLP as RewardMLP
from .reward_models.q_mlp import member_weights


# --------------------------------------------------------------------------------------------
def _keep_mask(dataset, env, terminate_on_end: bool, want_steps: bool = False):
    """iql.py:701-716 / 1236-1253 through the C scan (exact, O(N))."""
    n = int(dataset["rewards"].shape[0])
    term = np.ascontiguousarray(np.asarray(dataset["terminals"]).astype(np.uint8))
    tout = None
    if "timeouts" in dataset:
        tout = np.ascontiguousarray(np.asarray(dataset["timeouts"]).astype(np.uint8))
        max_steps = 0
    else:
        max_steps = int(env._max_episode_steps)
    keep = np.ones(max(n - 1, 0), dtype=np.uint8)
    steps = np.zeros(max(n - 1, 0), dtype=np.int32) if want_steps else None
    lib = _lib.load()
    _lib.check(lib.iqlb_keep_mask_scan(term.ctypes.data_as(C.c_void_p),
                                       tout.ctypes.data_as(C.c_void_p) if tout is not None else None, n, max_steps,
                                       int(bool(terminate_on_end)), keep.ctypes.data_as(C.c_void_p),
                                       steps.ctypes.data_as(C.c_void_p) if steps is not None else None),
               "iqlb_keep_mask_scan")
    return keep.astype(bool), steps


def _cuda_device(device) -> torch.device:
    dev = torch.device(device)
    if dev.type != "cuda" or not torch.cuda.is_available():
        raise RuntimeError(f"relabelling runs on the GPU only (device={device!r}); there is no CPU path")
    return dev


def cvar_n_tail(alpha: float, n_total: int) -> int:
    return max(1, int(np.floor((1.0 - alpha) * n_total)))  # iql.py:937,1148


RELABEL_PRECISIONS = {"auto": -1, "fp32": 0, "split_bf16": 1}  # include/iqlb.h IQLB_RELABEL_*


class PackedMembers:
    """S reward MLPs flattened into the [S, PM] fp32 device matrix `iqlb_relabel_mlp` reads (pack once, relabel
    many times: `ensemble_predict(pack_members(...), X)` skips the host-side flattening and the upload)."""

    def __init__(self, weights: torch.Tensor, d_in: int, widths: List[int]):
        self.weights, self.d_in, self.widths = weights, d_in, widths

    def __len__(self):
        return int(self.weights.shape[0])


def pack_members(members: Sequence[Sequence[Any]], device="cuda") -> PackedMembers:
    dev = _cuda_device(device)
    if len(members) == 0:
        raise ValueError("empty ensemble")
    first = [np.asarray(w.detach().cpu() if torch.is_tensor(w) else w) for w in members[0]]
    d_in = int(first[0].shape[0])
    widths = [int(first[2 * l].shape[1]) for l in range(len(first) // 2 - 1)]
    if int(first[-2].shape[1]) != 1:
        raise ValueError("reward MLPs must have a single output")
    flat = []
    for m in members:
        parts = [(w.detach().to("cpu", torch.float32).reshape(-1) if torch.is_tensor(w)
                  else torch.from_numpy(np.ascontiguousarray(np.asarray(w, dtype=np.float32)).reshape(-1))) for w in m]
        flat.append(torch.cat(parts))
    return PackedMembers(torch.stack(flat).to(dev), d_in, widths)


def ensemble_predict(members, X, reduce: str = "mean", alpha: float = 0.0,
                     activation: str = "relu", device="cuda", return_preds: bool = False,
                     return_selection: bool = False, precision: str = "auto", out: Optional[torch.Tensor] = None,
                     return_mean: bool = False):
    """Evaluate S reward MLPs on X [n, d_in] and reduce over members.

    members: S parameter lists [W0, b0, ..., Wout, bout] (numpy or torch; W stored (n_in, n_out) as
    in `RewardMLP.state_dict()` / the BNN `sampled_weights` tuples) or a `PackedMembers`.  reduce: "mean" or
    "cvar" (mean of the worst floor((1-alpha)*S) members, at least one).  precision: "fp32" (FFMA GEMMs),
    "split_bf16" (tcgen05, x_hi*W_hi + x_lo*W_hi + x_hi*W_lo; d_in <= 128, hidden [64|128] x <= 3) or "auto"
    (tensor cores where the shape allows).  X may be a row-strided view (stride(1) == 1) of a wider table.
    `out`: optional preallocated float32 CUDA tensor [n] to write into (repeated relabels without allocation).
    `return_mean`: also return the plain per-row mean over all members (same pass; no [S, n] matrix is materialised).
    Returns a float32 CUDA tensor [n] (plus, in this order, `preds` [S, n], the uint8 tail-membership mask [S, n] and
    the member mean [n] when asked for)."""
    dev = _cuda_device(device)
    if precision not in RELABEL_PRECISIONS:
        raise ValueError(f"precision must be one of {sorted(RELABEL_PRECISIONS)}, got {precision!r}")
    if reduce not in ("mean", "cvar"):
        raise ValueError(f"reduce must be 'mean' or 'cvar', got {reduce!r}")
    if reduce == "cvar" and not (0.0 <= alpha < 1.0):
        raise ValueError(f"alpha must be in [0, 1), got {alpha!r}")
    if activation not in _lib.ACTIVATIONS:
        raise ValueError(f"unknown activation {activation!r}")
    packed = members if isinstance(members, PackedMembers) else pack_members(members, dev)
    S, d_in, widths, weights = len(packed), packed.d_in, packed.widths, packed.weights
    X = torch.as_tensor(X, dtype=torch.float32)
    X = X.to(dev)
    if X.ndim != 2 or X.shape[1] != d_in:
        raise ValueError(f"X must be [n, {d_in}]")
    if X.shape[0] > 1 and (X.stride(1) != 1 or X.stride(0) < d_in):
        X = X.contiguous()
    n = X.shape[0]
    n_tail = S if reduce == "mean" else cvar_n_tail(alpha, S)
    if out is None:
        out = torch.empty(n, dtype=torch.float32, device=dev)
    elif tuple(out.shape) != (n,) or out.dtype != torch.float32 or out.device != X.device or not out.is_contiguous():
        raise ValueError(f"out must be a contiguous float32 tensor [{n}] on {X.device}")
    if n == 0:  # nothing to launch (an empty tensor has no device pointer)
        res = [out] + ([torch.empty((S, 0), dtype=torch.float32, device=dev)] if return_preds else []) + \
              ([torch.empty((S, 0), dtype=torch.uint8, device=dev)] if return_selection else []) + \
              ([torch.empty(0, dtype=torch.float32, device=dev)] if return_mean else [])
        return res[0] if len(res) == 1 else tuple(res)
    preds = torch.empty((S, n), dtype=torch.float32, device=dev) if return_preds else None
    sel = torch.empty((S, n), dtype=torch.uint8, device=dev) if return_selection else None
    mean = torch.empty(n, dtype=torch.float32, device=dev) if return_mean else None
    w_arr = (C.c_int32 * len(widths))(*widths)
    lib = _lib.load()
    ldx = X.stride(0) if n > 1 else d_in
    with _lib.device_ctx(dev):
        _lib.check(lib.iqlb_relabel_mlp_p(_lib.ptr(X), n, ldx, d_in, _lib.ptr(weights), S, w_arr, len(widths),
                                          _lib.ACTIVATIONS[activation], n_tail, RELABEL_PRECISIONS[precision],
                                          _lib.ptr(out), _lib.ptr(preds), _lib.ptr(sel), _lib.ptr(mean),
                                          _lib.stream_ptr(dev)),
                   "iqlb_relabel_mlp_p")
    res = [out]
    if return_preds:
        res.append(preds)
    if return_selection:
        res.append(sel)
    if return_mean:
        res.append(mean)
    return res[0] if len(res) == 1 else tuple(res)


def _activation_name(r_model) -> str:
    fn = getattr(r_model, "activation_name", None)
    if fn is not None:
        return fn
    f = r_model.activation_fn
    table = {torch.cos: "cos", torch.tanh: "tanh", torch.sin: "sin", torch.nn.functional.relu: "relu",
             torch.nn.functional.softplus: "softplus", torch.nn.functional.leaky_relu: "leaky_relu"}
    if f in table:
        return table[f]
    probe = f(torch.tensor([-1.0, 0.5]))  # the reference keeps lambdas for rbf/linear/swish
    cands = {"linear": torch.tensor([-1.0, 0.5]), "rbf": torch.exp(-torch.tensor([-1.0, 0.5]) ** 2),
             "swish": torch.tensor([-1.0, 0.5]) * torch.sigmoid(torch.tensor([-1.0, 0.5])),
             "sin": torch.sin(torch.tensor([-1.0, 0.5])), "cos": torch.cos(torch.tensor([-1.0, 0.5])),
             "softplus": torch.nn.functional.softplus(torch.tensor([-1.0, 0.5]))}
    for name, val in cands.items():
        if torch.allclose(probe, val, atol=1e-6):
            return name
    raise ValueError("cannot identify the reward model's activation function")


def _result(obs_all, act_all, keep, rewards, terminals):
    return {
        "observations": obs_all[:-1][keep],
        "actions": act_all[:-1][keep],
        "next_observations": obs_all[1:][keep],
        "rewards": rewards[keep],
        "terminals": terminals[:-1][keep],
    }


# --------------------------------------------------------------------------------------------
def qlearning_dataset_mr(env, r_model, dataset=None, terminate_on_end=False, **kwargs):
    """iql.py:691-732.  r_model: a reward MLP module (ours or the reference's optbnn MLP)."""
    if dataset is None:
        dataset = env.get_dataset(**kwargs)
    obs_all = dataset["observations"].astype(np.float32)
    act_all = dataset["actions"].astype(np.float32)
    keep, _ = _keep_mask(dataset, env, terminate_on_end)
    device = next(r_model.parameters()).device
    obs_act = np.concatenate([obs_all[:-1], act_all[:-1]], axis=1)
    rewards = ensemble_predict([member_weights(r_model)], obs_act, "mean", activation=_activation_name(r_model),
                               device=device).cpu().numpy()
    return _result(obs_all, act_all, keep, rewards, dataset["terminals"])


def empirical_cvar(samples: np.ndarray, alpha: float) -> float:
    """iql.py:735-763 (scalar diagnostic; host numpy exactly as the reference)."""
    if not (0.0 <= alpha < 1.0):
        raise ValueError(f"alpha must be in [0, 1), got {alpha!r}")
    sorted_samples = np.sort(samples)
    n_tail = max(1, int(np.floor((1.0 - alpha) * len(samples))))
    return float(sorted_samples[:n_tail].mean())


def cvar_check_indices(n_rows: int, n_checks: int = 50) -> np.ndarray:
    """The transitions `cvar_stability_check` looks at (iql.py:799-801: `default_rng(42).choice(N, min(n_checks, N))`)."""
    return np.random.default_rng(seed=42).choice(n_rows, size=min(n_checks, n_rows), replace=False)


def cvar_stability_check(all_preds, alpha: float, n_checks: int = 50, remedy: str = "Increase bnn_n_samples",
                         cols: Optional[np.ndarray] = None) -> float:
    """iql.py:766-827.  `all_preds` may be a CUDA tensor [S, N]; only the checked columns are copied.  `cols` [S, k]:
    the member predictions of the transitions `cvar_check_indices(N)` names, evaluated on their own (the relabel pass
    then never holds the [S, N] matrix the reference builds)."""
    if alpha == 0.0:
        print("[CVaR stability] alpha=0 (posterior mean); stability check skipped")
        return 0.0
    if cols is not None:
        cols = np.asarray(cols)
        S = cols.shape[0]
    else:
        S, N = all_preds.shape
        indices = cvar_check_indices(N, n_checks)
        if torch.is_tensor(all_preds):
            cols = all_preds[:, torch.as_tensor(indices, device=all_preds.device)].cpu().numpy()
        else:
            cols = all_preds[:, indices]
    ratios = []
    for j in range(cols.shape[1]):
        cvar_full = empirical_cvar(cols[:, j], alpha)
        cvar_half = empirical_cvar(cols[: S // 2, j], alpha)
        if abs(cvar_full) > 1e-8:
            ratios.append(abs(cvar_full - cvar_half) / abs(cvar_full))
    if not ratios:
        return float("nan")
    mean_ratio = float(np.mean(ratios))
    status = "OK" if mean_ratio < 0.05 else "WARN"
    print(f"[CVaR stability] mean relative diff = {mean_ratio:.3f} (target < 0.05) [{status}]")
    if mean_ratio > 0.05:
        min_s = int(np.ceil(30.0 / (1.0 - alpha)))
        warnings.warn(f"CVaR stability check: mean relative difference {mean_ratio:.3f} > 0.05. {remedy} "
                      f"(current S={S}). Recommended minimum for alpha={alpha}: S >= {min_s}.", RuntimeWarning)
    return mean_ratio


def _ensemble_rewards(members, obs_act, alpha, activation, device, tag, remedy, mean_label):
    """CVaR reward of every transition + the reference's two diagnostics, without the [S, N] prediction matrix of
    iql.py:982-1004 / 1162-1178 (3.9 GB at S = 981, N = 1M): the tail mean and the member mean come out of one pass over
    row chunks, and the stability check evaluates the ensemble once more on its <= 50 transitions only."""
    packed = members if isinstance(members, PackedMembers) else pack_members(members, _cuda_device(device))
    out, pm = ensemble_predict(packed, obs_act, "cvar", alpha, activation, device, return_mean=True)
    if alpha != 0.0:
        idx = cvar_check_indices(int(np.asarray(obs_act).shape[0]) if not torch.is_tensor(obs_act) else int(obs_act.shape[0]))
        x_chk = obs_act[torch.as_tensor(idx, device=obs_act.device)] if torch.is_tensor(obs_act) else np.asarray(obs_act)[idx]
        _, cols = ensemble_predict(packed, x_chk, "cvar", alpha, activation, device, return_preds=True)
        cvar_stability_check(None, alpha, remedy=remedy, cols=cols.cpu().numpy())
    else:
        cvar_stability_check(None, alpha, remedy=remedy, cols=np.zeros((len(packed), 0), np.float32))
    print(f"[{tag}] {mean_label}: {pm.mean().item():.4f} ± {pm.std(unbiased=False).item():.4f}")
    penalized = out.cpu().numpy()
    return penalized


def qlearning_dataset_bnn(env, reward_model_dir: str, alpha: float = 0.95, n_samples: int = 500, device: str = "cuda",
                          dataset=None, terminate_on_end: bool = False, **kwargs) -> Dict[str, np.ndarray]:
    """iql.py:830-1044."""
    if not (0.0 <= alpha < 1.0):
        raise ValueError(f"bnn_alpha must be in [0, 1), got {alpha!r}")
    if dataset is None:
        dataset = env.get_dataset(**kwargs)
    obs_all = dataset["observations"].astype(np.float32)
    act_all = dataset["actions"].astype(np.float32)
    keep, _ = _keep_mask(dataset, env, terminate_on_end)
    sampling_dir = os.path.join(reward_model_dir, "sampling_f")
    weight_files = sorted(_glob.glob(os.path.join(sampling_dir, "chain_*/sampled_weights/sampled_weights_0000000")))
    if not weight_files:
        raise FileNotFoundError(f"No BNN posterior weight files found under {sampling_dir}. Expected structure: "
                                "sampling_f/chain_*/sampled_weights/sampled_weights_0000000")
    all_weights: List = []
    for wf in weight_files:
        ckpt = torch.load(wf, map_location="cpu", weights_only=False)
        all_weights.extend(ckpt["sampled_weights"])
    if not all_weights:
        raise RuntimeError(f"BNN checkpoint at {reward_model_dir} contained no sampled weights.")
    available = len(all_weights)
    if n_samples > 0 and n_samples > available:
        warnings.warn(f"bnn_n_samples={n_samples} requested but only {available} posterior samples are available. "
                      f"Using all {available} samples. Add more chains / increase num_samples in the BNN training "
                      "config.", RuntimeWarning)
    if n_samples > 0 and n_samples < available:
        rng = np.random.default_rng(seed=0)
        idx = rng.choice(available, size=n_samples, replace=False)
        all_weights = [all_weights[i] for i in sorted(idx)]
    n_total = len(all_weights)
    n_tail = cvar_n_tail(alpha, n_total)
    min_recommended = int(np.ceil(30.0 / (1.0 - alpha)))
    if n_tail < 5:
        warnings.warn(f"CVaR tail has only {n_tail} sample(s) with alpha={alpha} and S={n_total}. CVaR estimates "
                      f"will be very noisy. Recommended minimum: bnn_n_samples >= {min_recommended}.", RuntimeWarning)
    print(f"[BNN/CVaR] S={n_total} samples from {len(weight_files)} chain(s), alpha={alpha}, n_tail={n_tail}")
    w0 = all_weights[0]
    print(f"[BNN/CVaR] Inferred architecture: input_dim={int(w0[0].shape[0])}, width={int(w0[0].shape[1])}, "
          f"depth={(len(w0) - 2) // 2}")
    obs_act = np.concatenate([obs_all[:-1], act_all[:-1]], axis=1)
    penalized_r = _ensemble_rewards(all_weights, obs_act, alpha, "relu", device, "BNN/CVaR", "Increase bnn_n_samples",
                                    "Posterior mean reward")
    label = "posterior mean reward" if alpha == 0.0 else f"CVaR reward (alpha={alpha})"
    print(f"[BNN/CVaR] {label}: {penalized_r.mean():.4f} ± {penalized_r.std():.4f}")
    if penalized_r.std() < 1e-6:
        warnings.warn("CVaR rewards have near-zero std — the penalty may have collapsed all rewards to the same "
                      "value. Consider a smaller alpha or larger bnn_n_samples.", RuntimeWarning)
    return _result(obs_all, act_all, keep, penalized_r, dataset["terminals"])


def _discover_mr_snapshots(reward_model_dir: str, burn_in: int = 0) -> List[str]:
    """iql.py:1047-1082."""
    epoch_re = re.compile(r"checkpoint_(\d+)\.pt$")
    found: List[Tuple[int, str]] = []
    for p in _glob.glob(os.path.join(reward_model_dir, "checkpoint_*.pt")):
        m = epoch_re.search(os.path.basename(p))
        if m is not None:
            found.append((int(m.group(1)), p))
    if not found:
        raise FileNotFoundError(f"No MR snapshots found in {reward_model_dir}. Expected per-epoch "
                                "checkpoint_{epoch}.pt files written by run_mr_training.py with checkpoints_path set.")
    kept = sorted((e, p) for e, p in found if e >= burn_in)
    if not kept:
        raise ValueError(f"mr_burn_in={burn_in} discarded all {len(found)} snapshot(s) in {reward_model_dir} "
                         f"(highest epoch present: {max(e for e, _ in found)}).")
    if len(kept) < len(found):
        print(f"[MR/CVaR] Burn-in {burn_in}: dropped {len(found) - len(kept)} snapshot(s) below epoch {burn_in}")
    return [p for _, p in kept]


def _strip_compile_prefix(state: Dict[str, Any]) -> Dict[str, Any]:
    """iql.py:1312-1323."""
    prefix = "_orig_mod."
    return {(k[len(prefix):] if k.startswith(prefix) else k): v for k, v in state.items()}


def _state_to_member(state: Dict[str, Any]) -> List[torch.Tensor]:
    """RewardMLP state_dict -> [W0, b0, W1, b1, ..., Wout, bout] (mlp.py:52-66 naming)."""
    out = [state["layers.0.W"], state["layers.0.b"]]
    i = 1
    while f"layers.linear_{i}.W" in state:
        out += [state[f"layers.linear_{i}.W"], state[f"layers.linear_{i}.b"]]
        i += 1
    out += [state["output_layer.W"], state["output_layer.b"]]
    return out


def _build_mlp_reward_model(state: Dict[str, Any], activations: str, device: str = "cuda"):
    """iql.py:1326-1336."""
    input_dim = state["layers.0.W"].shape[0]
    hidden_dims = [state["layers.0.W"].shape[1]]
    i = 1
    while f"layers.linear_{i}.W" in state:
        hidden_dims.append(state[f"layers.linear_{i}.W"].shape[1])
        i += 1
    return RewardMLP(input_dim, 1, hidden_dims, activations).to(device)


def _mr_activations(model_dir: str) -> str:
    with open(os.path.join(model_dir, "config.yaml")) as f:
        cfg = yaml.safe_load(f)
    return cfg.get("activations", "relu")


def qlearning_dataset_mr_ensemble(env, reward_model_dir: str, alpha: float = 0.95, burn_in: int = 0,
                                  device: str = "cuda", dataset=None, terminate_on_end: bool = False,
                                  **kwargs) -> Dict[str, np.ndarray]:
    """iql.py:1085-1220."""
    if not (0.0 <= alpha < 1.0):
        raise ValueError(f"mr_alpha must be in [0, 1), got {alpha!r}")
    if dataset is None:
        dataset = env.get_dataset(**kwargs)
    obs_all = dataset["observations"].astype(np.float32)
    act_all = dataset["actions"].astype(np.float32)
    keep, _ = _keep_mask(dataset, env, terminate_on_end)
    ckpt_paths = _discover_mr_snapshots(reward_model_dir, burn_in)
    n_total = len(ckpt_paths)
    n_tail = cvar_n_tail(alpha, n_total)
    if n_tail < 5 and alpha > 0.0:
        warnings.warn(f"CVaR tail has only {n_tail} snapshot(s) with alpha={alpha} and S={n_total}. The reward is "
                      "close to a per-transition worst-case over the ensemble rather than a tail average. Lower "
                      "mr_alpha or train the reward model with more eval epochs for a finer tail.", RuntimeWarning)
    print(f"[MR/CVaR] S={n_total} snapshot(s), alpha={alpha}, n_tail={n_tail}")
    activations = _mr_activations(reward_model_dir)
    members = []
    for p in ckpt_paths:
        ckpt = torch.load(p, map_location="cpu", weights_only=False)
        members.append(_state_to_member(_strip_compile_prefix(ckpt["net"])))
    obs_act = np.concatenate([obs_all[:-1], act_all[:-1]], axis=1)
    penalized_r = _ensemble_rewards(members, obs_act, alpha, activations, device, "MR/CVaR",
                                    "Lower mr_alpha, or retrain the reward model with more eval epochs to grow the "
                                    "ensemble", "Ensemble mean reward")
    label = "ensemble mean reward" if alpha == 0.0 else f"CVaR reward (alpha={alpha})"
    print(f"[MR/CVaR] {label}: {penalized_r.mean():.4f} ± {penalized_r.std():.4f}")
    if penalized_r.std() < 1e-6:
        warnings.warn("CVaR rewards have near-zero std — the penalty may have collapsed all rewards to the same "
                      "value. Consider a smaller mr_alpha.", RuntimeWarning)
    return _result(obs_all, act_all, keep, penalized_r, dataset["terminals"])


def pt_windows_from_steps(ep_steps: np.ndarray, query_length: int):
    """Window descriptors the reference builds implicitly (iql.py:1275-1292; SURVEY F4: the episode
    step is used as an ABSOLUTE row index).  Returns (win_start, seq_len) int32."""
    ep = ep_steps.astype(np.int64)
    full = ep >= query_length
    start = np.where(full, ep - query_length + 1, 0).astype(np.int32)
    seq = np.where(full, query_length, ep + 1).astype(np.int32)
    return start, seq


def pt_predict(r_model, obs_all: np.ndarray, act_all: np.ndarray, win_start: np.ndarray, seq_len: np.ndarray,
               query_length: int, device=None, dedupe: bool = True) -> torch.Tensor:
    """value[:, 0, -1, 0] of the preference transformer for every window (CUDA tensor [n])."""
    dev = _cuda_device(device if device is not None else next(r_model.parameters()).device)
    desc = pt_desc_from_model(r_model, query_length)
    weights = packed_pt_weights(r_model, desc, dev)
    obs = torch.as_tensor(obs_all, dtype=torch.float32).to(dev).contiguous()
    act = torch.as_tensor(act_all, dtype=torch.float32).to(dev).contiguous()
    # identical (start, length) windows give identical rewards: evaluate each distinct window once.
    # (With the reference's episode-step-as-row-index behaviour, SURVEY F4, a 1M-row dataset has at
    # most max_episode_steps distinct windows.)
    pairs = np.stack([np.asarray(win_start, dtype=np.int64), np.asarray(seq_len, dtype=np.int64)], axis=1)
    inverse = None
    if dedupe and pairs.shape[0] > 1:
        uniq, inverse = np.unique(pairs, axis=0, return_inverse=True)
        if uniq.shape[0] < pairs.shape[0]:
            win_start, seq_len = uniq[:, 0], uniq[:, 1]
        else:
            inverse = None
    ws = torch.as_tensor(np.asarray(win_start), dtype=torch.int32).to(dev).contiguous()
    sl = torch.as_tensor(np.asarray(seq_len), dtype=torch.int32).to(dev).contiguous()
    n = ws.shape[0]
    if n and int((ws.long() + sl.long()).max().item()) > obs.shape[0]:
        raise ValueError("a window reaches past the end of the dataset")
    out = torch.empty(n, dtype=torch.float32, device=dev)
    lib = _lib.load()
    with _lib.device_ctx(dev):
        _lib.check(lib.iqlb_relabel_pt(C.byref(desc), _lib.ptr(weights), _lib.ptr(obs), _lib.ptr(act), obs.shape[0],
                                       _lib.ptr(ws), _lib.ptr(sl), n, _lib.ptr(out), _lib.stream_ptr(dev)),
                   "iqlb_relabel_pt")
    if inverse is not None:
        out = out[torch.as_tensor(inverse.reshape(-1), dtype=torch.int64, device=dev)]
    return out


def qlearning_dataset_pt(env, r_model, query_length=100, dataset=None, terminate_on_end=False, **kwargs):
    """iql.py:1223-1309."""
    if dataset is None:
        dataset = env.get_dataset(**kwargs)
    obs_all = dataset["observations"].astype(np.float32)
    act_all = dataset["actions"].astype(np.float32)
    keep, ep_steps = _keep_mask(dataset, env, terminate_on_end, want_steps=True)
    start, seq = pt_windows_from_steps(ep_steps, query_length)
    rewards = pt_predict(r_model, obs_all, act_all, start, seq, query_length).cpu().numpy()
    return _result(obs_all, act_all, keep, rewards, dataset["terminals"])


def load_mlp_reward_model(model_dir: str, device: str = "cuda"):
    """iql.py:1345-1353."""
    ckpt = torch.load(os.path.join(model_dir, "best_model.pt"), map_location=device, weights_only=False)
    state = _strip_compile_prefix(ckpt["net"])
    model = _build_mlp_reward_model(state, _mr_activations(model_dir), device)
    model.load_state_dict(state)
    model.eval()
    return model


def load_pt_reward_model(model_dir: str, device: str = "cuda"):
    """iql.py:1356-1390 (all dimensions inferred from the state-dict shapes)."""
    with open(os.path.join(model_dir, "config.yaml")) as f:
        cfg = yaml.safe_load(f)
    ckpt = torch.load(os.path.join(model_dir, "best_model.pt"), map_location=device, weights_only=False)
    state = _strip_compile_prefix(ckpt["net"])
    embd_dim = state["state_linear.weight"].shape[0]
    num_layers = 0
    while f"gpt.layers.{num_layers}.layer_norm_0.weight" in state:
        num_layers += 1
    model = RewardPT(
        state_dim=state["state_linear.weight"].shape[1], action_dim=state["action_linear.weight"].shape[1],
        max_episode_steps=state["timestep_embed.weight"].shape[0] - 1, embd_dim=embd_dim,
        pref_attn_embd_dim=(state["pref_linear.weight"].shape[0] - 1) // 2, num_heads=cfg.get("num_heads", 4),
        attn_dropout=cfg.get("attn_dropout", 0.1), resid_dropout=cfg.get("resid_dropout", 0.1),
        intermediate_dim=cfg.get("intermediate_dim") or (4 * embd_dim), num_layers=num_layers,
        embd_dropout=cfg.get("embd_dropout", 0.1), max_pos=state["gpt.layers.0.attention.causal_bias"].shape[2],
        eps=cfg.get("model_eps", 1e-5)).to(device)
    model.load_state_dict(state)
    model.eval()
    return model
