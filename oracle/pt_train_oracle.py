"""CPU oracle of the preference-transformer training step (TEST INFRASTRUCTURE, not product).

A functional restatement, in plain torch on the CPU, of

* `PT.forward` in training mode (`gp_reward-priors/optbnn/bnn/nets/pref_trans.py:269-325`, `GPT2Model / GPT2Block /
  GPT2SelfAttention / GPT2MLP` `:69-227`): the three dropout sites (`embd_dropout` `:217`, `attn_dropout` inside SDPA
  `:134-138`, `resid_dropout` after the attention projection `:142` and after the MLP `:86`) take EXPLICIT masks
  (0 or 1/(1-p)), because the reference draws them from torch's generator, which the CUDA path cannot mirror;
* `PTTrainer.train / .evaluation` (`optbnn/training/training.py:308-416`): one forward over both segments, Bradley-Terry
  pooling of `weighted_sum` (`bt_pool_logit`, `optbnn/utils/util.py:356-364`), soft-label cross entropy, accuracy;
* `torch.optim.Adam(net.parameters(), lr)` (`scripts_pt/run_pt_training.py:137`).

Gradients come from torch.autograd on this restatement (as in oracle/iql_oracle.py).  Pinned to fixtures written by
the unmodified reference `PTTrainer` with all dropouts at 0 (`oracle/gen_golden_pt.py` -> `tests/golden/pt_train.npz`,
`tests/test_pt_oracle_golden.py`).
"""
import math
from typing import Dict, List, Optional

import numpy as np
import torch
import torch.nn.functional as F

SITE_EMBD = 0


def site_attn(layer: int) -> int:
    return 1 + 4 * layer


def site_resid_attn(layer: int) -> int:
    return 2 + 4 * layer


def site_resid_mlp(layer: int) -> int:
    return 3 + 4 * layer


def n_layers(p: Dict[str, torch.Tensor]) -> int:
    n = 0
    while f"gpt.layers.{n}.layer_norm_0.weight" in p:
        n += 1
    return n


def pt_forward(p: Dict[str, torch.Tensor], states, actions, timesteps, attn_mask, num_heads: int, eps: float,
               masks: Optional[Dict[int, torch.Tensor]] = None):
    """Returns (weighted_sum [n, T], value [n, T]).  `masks[site]`: multiplicative dropout mask of that site (shape of
    the tensor it multiplies: [n, 2T, E], or [n, H, 2T, 2T] for the attention probabilities); missing = no dropout."""
    masks = masks or {}
    n, t, _ = states.shape
    e = p["state_linear.weight"].shape[0]
    temb = p["timestep_embed.weight"][timesteps.long()]
    es = F.linear(states, p["state_linear.weight"], p["state_linear.bias"]) + temb
    ea = F.linear(actions, p["action_linear.weight"], p["action_linear.bias"]) + temb
    x = torch.stack([es, ea], dim=1).permute(0, 2, 1, 3).reshape(n, 2 * t, e)  # pref_trans.py:280-284
    x = F.layer_norm(x, (e,), p["stacked_layer_norm.weight"], p["stacked_layer_norm.bias"], eps)
    if SITE_EMBD in masks:
        x = x * masks[SITE_EMBD]
    m2 = torch.stack([attn_mask, attn_mask], dim=1).permute(0, 2, 1).reshape(n, 2 * t)  # :288-292
    pad = ((1.0 - m2) * -10000.0)[:, None, None, :]
    causal = torch.triu(torch.full((2 * t, 2 * t), float("-inf")), diagonal=1)[None, None]
    dh = e // num_heads
    for l in range(n_layers(p)):
        pre = f"gpt.layers.{l}."
        h = F.layer_norm(x, (e,), p[pre + "layer_norm_0.weight"], p[pre + "layer_norm_0.bias"], eps)
        qkv = F.linear(h, p[pre + "attention.in_linear.weight"], p[pre + "attention.in_linear.bias"])
        q, k, v = torch.chunk(qkv, 3, dim=2)
        sh = lambda z: z.reshape(n, 2 * t, num_heads, dh).transpose(1, 2)
        q, k, v = sh(q), sh(k), sh(v)
        s = q @ k.transpose(-1, -2) * (1.0 / math.sqrt(dh)) + causal + pad
        w = torch.softmax(s, dim=-1)
        if site_attn(l) in masks:
            w = w * masks[site_attn(l)]
        o = (w @ v).transpose(1, 2).reshape(n, 2 * t, e)
        o = F.linear(o, p[pre + "attention.out_linear.weight"], p[pre + "attention.out_linear.bias"])
        if site_resid_attn(l) in masks:
            o = o * masks[site_resid_attn(l)]
        x = x + o
        h = F.layer_norm(x, (e,), p[pre + "layer_norm_1.weight"], p[pre + "layer_norm_1.bias"], eps)
        h = torch.relu(F.linear(h, p[pre + "mlp.in_linear.weight"], p[pre + "mlp.in_linear.bias"]))
        h = F.linear(h, p[pre + "mlp.out_linear.weight"], p[pre + "mlp.out_linear.bias"])
        if site_resid_mlp(l) in masks:
            h = h * masks[site_resid_mlp(l)]
        x = x + h
    x = F.layer_norm(x, (e,), p["gpt.layer_norm.weight"], p["gpt.layer_norm.bias"], eps)
    hid = x.reshape(n, t, 2, e)[:, :, 1]  # the action tokens (:300-301)
    y = F.linear(hid, p["pref_linear.weight"], p["pref_linear.bias"])
    pd = (y.shape[-1] - 1) // 2
    q, k, v = y[..., :pd], y[..., pd:2 * pd], y[..., 2 * pd]
    s = q @ k.transpose(-1, -2) * (1.0 / math.sqrt(pd)) + ((1.0 - attn_mask) * -10000.0)[:, None, :]
    w = torch.softmax(s, dim=-1)
    return torch.einsum("btj,bj->bt", w, v), v


def bt_pool_logit(pred, mask, mode: str):
    """optbnn/utils/util.py:356-364."""
    s = torch.nansum(pred, dim=1)
    if mode == "sum":
        return s
    return s / torch.nansum(mask, dim=1).clamp(min=1.0)


def pt_loss(p, batch, num_heads, eps, bt_pool="mean", masks=None):
    """PTTrainer.train's forward (training.py:333-352): (loss, accuracy)."""
    s1, a1, t1, m1, s2, a2, t2, m2, labels = batch
    b = s1.shape[0]
    ws, _ = pt_forward(p, torch.cat([s1, s2]), torch.cat([a1, a2]), torch.cat([t1, t2]), torch.cat([m1, m2]), num_heads, eps,
                       masks)
    l1 = bt_pool_logit(ws[:b] * m1, m1, bt_pool)
    l2 = bt_pool_logit(ws[b:] * m2, m2, bt_pool)
    logits = torch.stack([l1, l2], dim=1)
    loss = F.cross_entropy(logits, labels)
    acc = (logits.argmax(1) == labels.argmax(1)).float().mean()
    return loss, acc


class PTTrainerOracle:
    """K training steps on a parameter dict (numpy in, cloned)."""

    def __init__(self, params: Dict[str, np.ndarray], num_heads: int, eps: float, lr: float = 3e-4, bt_pool: str = "mean"):
        self.p = {k: torch.tensor(np.asarray(v), dtype=torch.float32, requires_grad=True) for k, v in params.items()}
        self.num_heads, self.eps, self.lr, self.bt_pool = num_heads, eps, lr, bt_pool
        self.m = {k: torch.zeros_like(v) for k, v in self.p.items()}
        self.v = {k: torch.zeros_like(v) for k, v in self.p.items()}
        self.step = 0

    def _batch(self, batch):
        return [torch.as_tensor(np.asarray(x)) for x in batch]

    def loss_and_grads(self, batch, masks=None):
        loss, acc = pt_loss(self.p, self._batch(batch), self.num_heads, self.eps, self.bt_pool, masks)
        keys = list(self.p)
        grads = torch.autograd.grad(loss, [self.p[k] for k in keys], allow_unused=True)
        g = {k: (torch.zeros_like(self.p[k]) if gr is None else gr) for k, gr in zip(keys, grads)}
        return float(loss), float(acc), g

    def train(self, batch, masks=None) -> Dict[str, float]:
        loss, acc, g = self.loss_and_grads(batch, masks)
        self.step += 1
        b1, b2, eps = 0.9, 0.999, 1e-8
        bc1, bc2 = 1 - b1 ** self.step, 1 - b2 ** self.step
        with torch.no_grad():
            for k, prm in self.p.items():
                self.m[k].lerp_(g[k], 1 - b1)
                self.v[k].mul_(b2).addcmul_(g[k], g[k], value=1 - b2)
                prm.addcdiv_(self.m[k], (self.v[k].sqrt() / math.sqrt(bc2)).add_(eps), value=-self.lr / bc1)
        return {"training_loss": loss, "training_acc": acc}

    def evaluation(self, batch) -> Dict[str, float]:
        with torch.no_grad():
            loss, acc = pt_loss(self.p, self._batch(batch), self.num_heads, self.eps, self.bt_pool)
        return {"eval_loss": float(loss), "eval_acc": float(acc)}

    def weights(self) -> Dict[str, np.ndarray]:
        return {k: v.detach().numpy().copy() for k, v in self.p.items()}


def synth_pt_batch(rng, B, T, S, A, ragged=True):
    """(states, actions, timesteps, attn_mask) x 2 + one-hot / soft labels, left-padded like the relabel windows."""
    out = []
    for _ in range(2):
        s = rng.standard_normal((B, T, S)).astype(np.float32)
        a = rng.uniform(-1, 1, (B, T, A)).astype(np.float32)
        ts = np.tile(np.arange(T, dtype=np.int64), (B, 1))
        m = np.ones((B, T), np.float32)
        if ragged:
            for j in range(0, B, 3):
                padn = int(rng.integers(0, T))
                s[j, :padn] = 0; a[j, :padn] = 0; m[j, :padn] = 0
                ts[j, :padn] = 0; ts[j, padn:] = np.arange(T - padn)
        out += [s, a, ts, m]
    y = (rng.uniform(size=B) < 0.5).astype(np.float32)
    labels = np.stack([y, 1 - y], axis=1)
    labels[0] = [0.5, 0.5]
    out.append(labels.astype(np.float32))
    return out


def philox_dropout_mask(seed: int, adam_step: int, site: int, shape, p: float) -> np.ndarray:
    """The keep mask the CUDA step applies (include/iqlb.h, csrc/pt_train.cu): element e (row-major) of `site` at Adam
    step k is kept iff word (e & 3) of Philox4x32-10(key = seed, counter = {lo32(k), hi32(k), lo32(e >> 2),
    (site << 8) | hi32(e >> 2)}) >= floor(p * 2^32); kept elements carry 1 / (1 - p) (float32 arithmetic)."""
    from oracle.philox_oracle import philox4x32_10
    numel = int(np.prod(shape))
    if p <= 0:
        return np.ones(shape, np.float32)
    q = np.arange((numel + 3) // 4, dtype=np.uint64)
    ctr = np.stack([np.full_like(q, adam_step & 0xFFFFFFFF), np.full_like(q, adam_step >> 32), q & np.uint64(0xFFFFFFFF),
                    (np.uint64(site) << np.uint64(8)) | (q >> np.uint64(32))], axis=-1)
    key = np.stack([np.full_like(q, seed & 0xFFFFFFFF), np.full_like(q, (seed >> 32) & 0xFFFFFFFF)], axis=-1)
    words = philox4x32_10(ctr, key).reshape(-1)[:numel]
    thresh = min(int(np.float64(np.float32(p)) * 4294967296.0), 0xFFFFFFFF)
    scale = np.float32(1.0) / (np.float32(1.0) - np.float32(p))
    return np.where(words >= max(thresh, 1), scale, np.float32(0)).astype(np.float32).reshape(shape)


def dropout_masks(seed: int, adam_step: int, n: int, T: int, E: int, H: int, L: int, p_embd: float, p_attn: float,
                  p_resid_attn: float, p_resid_mlp: float) -> Dict[int, torch.Tensor]:
    """All sites of one training step, keyed like `pt_forward`'s `masks`."""
    m = {}
    if p_embd > 0:
        m[SITE_EMBD] = torch.from_numpy(philox_dropout_mask(seed, adam_step, SITE_EMBD, (n, 2 * T, E), p_embd))
    for l in range(L):
        if p_attn > 0:
            m[site_attn(l)] = torch.from_numpy(philox_dropout_mask(seed, adam_step, site_attn(l), (n, H, 2 * T, 2 * T), p_attn))
        if p_resid_attn > 0:
            m[site_resid_attn(l)] = torch.from_numpy(
                philox_dropout_mask(seed, adam_step, site_resid_attn(l), (n, 2 * T, E), p_resid_attn))
        if p_resid_mlp > 0:
            m[site_resid_mlp(l)] = torch.from_numpy(
                philox_dropout_mask(seed, adam_step, site_resid_mlp(l), (n, 2 * T, E), p_resid_mlp))
    return m
