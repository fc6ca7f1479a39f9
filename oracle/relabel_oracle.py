"""CPU oracle for the reward-model relabelling pass (TEST INFRASTRUCTURE, not product).

numpy restatement of:

* the keep-mask / episode-step scan shared by all four `qlearning_dataset_*`
  (`/root/reference/algorithms/offline/iql.py:701-716`, `:1236-1253`),
* `RewardMLP.forward` (`gp_reward-priors/optbnn/bnn/nets/mlp.py:73-93`) with the
  scaled-variance `Linear` (`optbnn/bnn/layers/linear.py:34-47`: X @ (W/sqrt(n_in)) + b,
  W stored (n_in, n_out)) and the activations of `optbnn/bnn/activation_fns.py:10-24`,
* the ensemble CVaR reduction (`iql.py:937,1009-1011`, `:1148,1185-1187`) and
  `empirical_cvar` (`iql.py:735-763`),
* the preference-transformer forward as the relabel uses it
  (`optbnn/bnn/nets/pref_trans.py:269-325`, GPT2 block `:118-183`), returning
  `value[:, 0, -1, 0]`, and the window construction of `qlearning_dataset_pt`
  (`iql.py:1259-1301`, including the episode-step-as-row-index behaviour),
* `return_reward_range` / `modify_reward` (`iql.py:344-401`).

Parity pinning: the reference has no tests for this path; pinned against outputs of the
reference itself (`oracle/gen_golden.py` -> `tests/golden/relabel_*.npz`).
"""
import math
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

f32 = np.float32


# ---------------------------------------------------------------------------- scan
def keep_mask_scan(terminals: np.ndarray, timeouts: Optional[np.ndarray], max_episode_steps: int,
                   terminate_on_end: bool = False) -> Tuple[np.ndarray, np.ndarray]:
    """iql.py:701-716 / 1236-1253.  Returns (keep[N-1] bool, ep_steps[N-1] int64)."""
    n = terminals.shape[0]
    keep = np.ones(n - 1, dtype=bool)
    ep_steps = np.zeros(n - 1, dtype=np.int64)
    ep = 0
    for i in range(n - 1):
        ep_steps[i] = ep
        done = bool(terminals[i])
        final = bool(timeouts[i]) if timeouts is not None else ep == max_episode_steps - 1
        if (not terminate_on_end) and final:
            keep[i] = False
            ep = 0
            continue
        if done or final:
            ep = 0
        ep += 1
    return keep, ep_steps


# ---------------------------------------------------------------------------- reward MLP
def _act(name: str, x: np.ndarray) -> np.ndarray:
    if name == "relu":
        return np.maximum(x, f32(0))
    if name == "tanh":
        return np.tanh(x)
    if name == "cos":
        return np.cos(x)
    if name == "sin":
        return np.sin(x)
    if name == "linear":
        return x
    if name == "rbf":
        return np.exp(-x * x)
    if name == "swish":
        return x * (f32(1) / (f32(1) + np.exp(-x)))
    if name == "softplus":  # F.softplus beta=1 threshold=20
        return np.where(x > 20, x, np.log1p(np.exp(np.minimum(x, f32(20)))))
    if name == "leaky_relu":
        return np.where(x > 0, x, f32(0.01) * x)
    raise ValueError(name)


def reward_mlp_forward(x: np.ndarray, weights: Sequence[np.ndarray], activation: str = "relu") -> np.ndarray:
    """weights = [W0, b0, W1, b1, ..., W_out, b_out] with W (n_in, n_out).  Returns [n, out]."""
    x = np.asarray(x, dtype=f32)
    n_layers = len(weights) // 2
    for l in range(n_layers):
        w = np.asarray(weights[2 * l], dtype=f32)
        b = np.asarray(weights[2 * l + 1], dtype=f32)
        ws = w / f32(math.sqrt(w.shape[0]))
        x = x @ ws + b
        if l < n_layers - 1:
            x = _act(activation, x).astype(f32)
    return x


def cvar_n_tail(alpha: float, s: int) -> int:
    return max(1, int(np.floor((1.0 - alpha) * s)))


def cvar_reduce(all_preds: np.ndarray, alpha: float) -> np.ndarray:
    """(S, N) -> (N,): mean of the n_tail smallest per column (iql.py:1009-1011)."""
    s = all_preds.shape[0]
    n_tail = cvar_n_tail(alpha, s)
    srt = np.sort(all_preds, axis=0)
    return srt[:n_tail].mean(axis=0).astype(f32)


def cvar_selection(all_preds: np.ndarray, alpha: float) -> np.ndarray:
    """Boolean (S, N) membership of the n_tail smallest per column (stable order on ties:
    lower member index first).  The set the relabel kernel must reproduce bit-exactly."""
    s, n = all_preds.shape
    n_tail = cvar_n_tail(alpha, s)
    order = np.argsort(all_preds, axis=0, kind="stable")
    sel = np.zeros((s, n), dtype=bool)
    np.put_along_axis(sel, order[:n_tail], True, axis=0)
    return sel


def ensemble_relabel(x: np.ndarray, members: Sequence[Sequence[np.ndarray]], activation: str,
                     alpha: float) -> np.ndarray:
    preds = np.stack([reward_mlp_forward(x, w, activation)[:, 0] for w in members], axis=0)
    return cvar_reduce(preds, alpha)


# ---------------------------------------------------------------------------- PT
def _layer_norm(x, g, b, eps):
    mu = x.mean(-1, keepdims=True)
    var = ((x - mu) ** 2).mean(-1, keepdims=True)
    return ((x - mu) / np.sqrt(var + f32(eps)) * g + b).astype(f32)


def pt_forward_value(p: Dict[str, np.ndarray], states, actions, timesteps, attn_mask,
                     num_heads: int, eps: float, with_weighted_sum: bool = False):
    """pref_trans.py:269-325, eval mode.  p = state_dict as numpy.  Returns value [B, T]
    (and, with_weighted_sum, weighted_sum [B, T]: the preference-attention head :303-320)."""
    states = np.asarray(states, f32)
    actions = np.asarray(actions, f32)
    bsz, t, _ = states.shape
    e = p["state_linear.weight"].shape[0]
    temb = p["timestep_embed.weight"][timesteps]
    es = states @ p["state_linear.weight"].T + p["state_linear.bias"] + temb
    ea = actions @ p["action_linear.weight"].T + p["action_linear.bias"] + temb
    x = np.stack([es, ea], axis=2).reshape(bsz, 2 * t, e).astype(f32)
    x = _layer_norm(x, p["stacked_layer_norm.weight"], p["stacked_layer_norm.bias"], eps)
    m2 = np.repeat(np.asarray(attn_mask, f32), 2, axis=1)  # [B, 2T]
    pad = (f32(1) - m2) * f32(-10000.0)
    causal = np.triu(np.full((2 * t, 2 * t), -np.inf, dtype=f32), k=1)
    bias = causal[None, None] + pad[:, None, None, :]
    dh = e // num_heads
    layer = 0
    while f"gpt.layers.{layer}.layer_norm_0.weight" in p:
        pre = f"gpt.layers.{layer}."
        h = _layer_norm(x, p[pre + "layer_norm_0.weight"], p[pre + "layer_norm_0.bias"], eps)
        qkv = h @ p[pre + "attention.in_linear.weight"].T + p[pre + "attention.in_linear.bias"]
        q, k, v = np.split(qkv, 3, axis=-1)
        sh = lambda z: z.reshape(bsz, 2 * t, num_heads, dh).transpose(0, 2, 1, 3)
        q, k, v = sh(q), sh(k), sh(v)
        s = (q @ k.transpose(0, 1, 3, 2)) * f32(1.0 / math.sqrt(dh)) + bias
        s = s - s.max(-1, keepdims=True)
        w = np.exp(s)
        w = w / w.sum(-1, keepdims=True)
        o = (w @ v).transpose(0, 2, 1, 3).reshape(bsz, 2 * t, e)
        o = o @ p[pre + "attention.out_linear.weight"].T + p[pre + "attention.out_linear.bias"]
        x = (x + o).astype(f32)
        h = _layer_norm(x, p[pre + "layer_norm_1.weight"], p[pre + "layer_norm_1.bias"], eps)
        h = np.maximum(h @ p[pre + "mlp.in_linear.weight"].T + p[pre + "mlp.in_linear.bias"], 0)
        h = h @ p[pre + "mlp.out_linear.weight"].T + p[pre + "mlp.out_linear.bias"]
        x = (x + h).astype(f32)
        layer += 1
    x = _layer_norm(x, p["gpt.layer_norm.weight"], p["gpt.layer_norm.bias"], eps)
    hid = x.reshape(bsz, t, 2, e)[:, :, 1]  # action tokens
    wv = p["pref_linear.weight"][-1]
    bv = p["pref_linear.bias"][-1]
    value = (hid @ wv + bv).astype(f32)  # value [B, T]
    if not with_weighted_sum:
        return value
    # preference attention (pref_trans.py:303-320): one head, not causal, additive key-padding mask, scale 1/sqrt(P)
    y = (hid @ p["pref_linear.weight"].T + p["pref_linear.bias"]).astype(f32)  # [B, T, 2P+1]
    pd = (y.shape[-1] - 1) // 2
    q, k, v = y[..., :pd], y[..., pd:2 * pd], y[..., 2 * pd]
    s = (q @ k.transpose(0, 2, 1)) * f32(1.0 / math.sqrt(pd)) + ((f32(1) - np.asarray(attn_mask, f32)) * f32(-10000.0))[:, None, :]
    s = s - s.max(-1, keepdims=True)
    w = np.exp(s)
    w = w / w.sum(-1, keepdims=True)
    return value, np.einsum("btj,bj->bt", w, v).astype(f32)


def pt_windows(obs_all: np.ndarray, act_all: np.ndarray, ep_steps: np.ndarray, query_length: int):
    """iql.py:1264-1292 for an arbitrary slice: returns (sts, acts, ts, am)."""
    b = ep_steps.shape[0]
    s_dim, a_dim = obs_all.shape[1], act_all.shape[1]
    sts = np.zeros((b, query_length, s_dim), f32)
    acts = np.zeros((b, query_length, a_dim), f32)
    ts = np.zeros((b, query_length), np.int64)
    am = np.zeros((b, query_length), f32)
    tmpl = np.arange(query_length, dtype=np.int64)
    for j in range(b):
        ep = int(ep_steps[j])
        if ep >= query_length:
            rows = ep - query_length + 1 + tmpl
            sts[j] = obs_all[rows]
            acts[j] = act_all[rows]
            ts[j] = tmpl
            am[j] = 1.0
        else:
            seq = ep + 1
            padn = query_length - seq
            sts[j, padn:] = obs_all[:seq]
            acts[j, padn:] = act_all[:seq]
            ts[j, padn:] = tmpl[:seq]
            am[j, padn:] = 1.0
    return sts, acts, ts, am


def pt_relabel(p, obs_all, act_all, ep_steps, query_length, num_heads, eps, chunk=64) -> np.ndarray:
    out = np.zeros(ep_steps.shape[0], f32)
    for cs in range(0, ep_steps.shape[0], chunk):
        sl = slice(cs, min(cs + chunk, ep_steps.shape[0]))
        sts, acts, ts, am = pt_windows(obs_all, act_all, ep_steps[sl], query_length)
        out[sl] = pt_forward_value(p, sts, acts, ts, am, num_heads, eps)[:, -1]
    return out


# ---------------------------------------------------------------------------- rewards
def return_reward_range(rewards, terminals, max_episode_steps):
    """iql.py:344-360."""
    returns, lengths = [], []
    ep_ret, ep_len, i = 0.0, 0, 0
    trj = np.zeros(rewards.shape[0])
    for j in range(rewards.shape[0]):
        ep_ret += float(rewards[j])
        ep_len += 1
        trj[i:j + 1] = ep_len
        if terminals[j] or ep_len == max_episode_steps:
            i = j + 1
            returns.append(ep_ret)
            lengths.append(ep_len)
            ep_ret, ep_len = 0.0, 0
    return min(returns), max(returns), trj


def modify_reward(rewards: np.ndarray, terminals: np.ndarray, env_name: str, normalize_reward: int,
                  max_episode_steps: int = 1000) -> np.ndarray:
    """iql.py:363-401 on a copy (the in-place numpy op sequence is kept, dtype included)."""
    r = rewards.copy()
    if any(s in env_name for s in ("halfcheetah", "hopper", "walker2d")):
        mn, mx, _ = return_reward_range(r, terminals, max_episode_steps)
        r /= mx - mn
        r *= max_episode_steps
    elif "antmaze" in env_name:
        if normalize_reward == 1:
            r -= 1.0
        else:
            mn, mx, trj = return_reward_range(r, terminals, max_episode_steps)
            if normalize_reward in (4, 5):
                r -= mn
            elif normalize_reward not in (2, 3):
                r -= mn / trj
            r /= mx - mn
            r *= max_episode_steps
            if normalize_reward in (3, 5) or normalize_reward >= 7 or normalize_reward < 1:
                r -= 1.0
    return r
