"""Preference transformer.  Mirrors `optbnn/bnn/nets/pref_trans.py:230-325` (the PyTorch twin that
`algorithms/offline/iql.py:33` imports as RewardPT) and the call surface of the JAX `PT`
(`reward_models/pref_transformer.py:210-277`: `(states, actions, timesteps, attn_mask)`).

Parameter container with the reference's state_dict keys; `forward` runs the CUDA forward and
returns `({"weighted_sum": [B,T,1], "value": [B,1,T,1]}, [None]*(L+1))` exactly like the reference (whose SDPA
returns no attention weights, pref_trans.py:321).  The relabel pass reads only `value[:, 0, -1, 0]` (iql.py:1301);
`weighted_sum` (the preference-attention head, :303-325) is what the Bradley-Terry loss of `PTTrainer` pools
(optbnn/training/training.py:348).
"""
import ctypes as C
from typing import Dict

import torch
import torch.nn as nn

from .. import _lib


class _SelfAttention(nn.Module):
    def __init__(self, embd_dim, num_heads, max_pos):
        super().__init__()
        self.num_heads = num_heads
        self.head_dim = embd_dim // num_heads
        self.in_linear = nn.Linear(embd_dim, 3 * embd_dim)
        self.out_linear = nn.Linear(embd_dim, embd_dim)
        self.register_buffer("causal_bias", torch.triu(torch.full((max_pos, max_pos), float("-inf")), diagonal=1)
                             .view(1, 1, max_pos, max_pos))


class _MLP(nn.Module):
    def __init__(self, embd_dim, intermediate_dim):
        super().__init__()
        self.in_linear = nn.Linear(embd_dim, intermediate_dim)
        self.out_linear = nn.Linear(intermediate_dim, embd_dim)


class _Block(nn.Module):
    def __init__(self, embd_dim, num_heads, intermediate_dim, max_pos, eps):
        super().__init__()
        self.layer_norm_0 = nn.LayerNorm(embd_dim, eps)
        self.attention = _SelfAttention(embd_dim, num_heads, max_pos)
        self.layer_norm_1 = nn.LayerNorm(embd_dim, eps)
        self.mlp = _MLP(embd_dim, intermediate_dim)


class _GPT2(nn.Module):
    def __init__(self, embd_dim, num_heads, intermediate_dim, num_layers, max_pos, eps):
        super().__init__()
        self.layers = nn.ModuleList([_Block(embd_dim, num_heads, intermediate_dim, max_pos, eps)
                                     for _ in range(num_layers)])
        self.layer_norm = nn.LayerNorm(embd_dim, eps)


class PT(nn.Module):
    def __init__(self, state_dim: int = 16, action_dim: int = 3, max_episode_steps: int = 1219, embd_dim: int = 64,
                 pref_attn_embd_dim: int = 64, num_heads: int = 4, attn_dropout: float = 0.1,
                 resid_dropout: float = 0.1, intermediate_dim: int = 256, num_layers: int = 1,
                 embd_dropout: float = 0.1, max_pos: int = 1024, eps: float = 1e-05):
        super().__init__()
        if embd_dim % num_heads:
            raise ValueError("embd_dim must be divisible by num_heads")
        self.embd_dim, self.pref_attn_embd_dim, self.num_heads = embd_dim, pref_attn_embd_dim, num_heads
        self.intermediate_dim, self.num_layers, self.eps = intermediate_dim, num_layers, eps
        # dropout probabilities of training mode (iqlpref_b200.pt_training.PTTrainer; forward() here is eval mode).  The
        # reference's GPT2Model does not hand `resid_dropout` to its blocks (pref_trans.py:203-213), so both residual sites
        # of every block run at GPT2Block's default 0.1 whatever the argument says; mirrored.
        self.embd_dropout_p, self.attn_dropout_p, self.resid_dropout_p = float(embd_dropout), float(attn_dropout), 0.1
        self.state_linear = nn.Linear(state_dim, embd_dim)
        self.action_linear = nn.Linear(action_dim, embd_dim)
        self.timestep_embed = nn.Embedding(max_episode_steps + 1, embd_dim)
        self.stacked_layer_norm = nn.LayerNorm(embd_dim, eps)
        self.gpt = _GPT2(embd_dim, num_heads, intermediate_dim, num_layers, max_pos, eps)
        self.pref_linear = nn.Linear(embd_dim, 2 * pref_attn_embd_dim + 1)

    def forward(self, states, actions, timesteps, attn_mask):
        """Explicit windows [B,T,S], [B,T,A], [B,T] int64, [B,T] float -> value [B,1,T,1]."""
        dev = states.device
        if dev.type != "cuda":
            raise RuntimeError("PT.forward runs on the GPU only (no CPU path)")
        b, t, _ = states.shape
        desc = pt_desc_from_model(self, t)
        weights = packed_pt_weights(self, desc, dev)
        st = states.to(torch.float32).contiguous()
        ac = actions.to(torch.float32).contiguous()
        ts = timesteps.to(torch.int32).contiguous()
        am = attn_mask.to(torch.float32).contiguous()
        out = torch.empty((b, t), dtype=torch.float32, device=dev)
        wsum = torch.empty((b, t), dtype=torch.float32, device=dev)
        lib = _lib.load()
        with _lib.device_ctx(dev):
            _lib.check(lib.iqlb_pt_forward(C.byref(desc), _lib.ptr(weights), _lib.ptr(st), _lib.ptr(ac), _lib.ptr(ts),
                                           _lib.ptr(am), b, _lib.ptr(out), _lib.ptr(wsum), _lib.stream_ptr(dev)),
                       "iqlb_pt_forward")
        return {"weighted_sum": wsum.view(b, t, 1), "value": out.view(b, 1, t, 1)}, [None] * (self.num_layers + 1)


def pt_desc_from_model(model, query_len: int) -> "_lib.PTDesc":
    """Works for our PT and for the reference's optbnn PT (same attribute / parameter names)."""
    sd = model.state_dict()
    e = sd["state_linear.weight"].shape[0]
    layers = 0
    while f"gpt.layers.{layers}.layer_norm_0.weight" in sd:
        layers += 1
    heads = model.gpt.layers[0].attention.num_heads
    return _lib.PTDesc(sd["state_linear.weight"].shape[1], sd["action_linear.weight"].shape[1], e, heads,
                       sd["gpt.layers.0.mlp.in_linear.weight"].shape[0], layers,
                       (sd["pref_linear.weight"].shape[0] - 1) // 2, sd["timestep_embed.weight"].shape[0], query_len,
                       float(model.stacked_layer_norm.eps))


def packed_pt_weights(model, desc, device) -> torch.Tensor:
    """`pack_pt_weights(model, desc)` on `device`, cached on the module until one of its parameters changes (the relabel
    calls the model thousands of times with the same weights; packing goes through the host)."""
    key = (str(device), desc.query_len) + tuple((p.data_ptr(), p._version) for p in model.parameters())
    cached = getattr(model, "_iqlb_packed", None)
    if cached is not None and cached[0] == key:
        return cached[1]
    w = pack_pt_weights(model, desc).to(device)
    try:
        object.__setattr__(model, "_iqlb_packed", (key, w))
    except Exception:  # noqa: BLE001
        pass
    return w


def pack_pt_weights(model, desc) -> torch.Tensor:
    """Flat fp32 weights in the order csrc/relabel_pt.cu expects (see PtLayout there)."""
    sd: Dict[str, torch.Tensor] = {k: v.detach().to("cpu", torch.float32) for k, v in model.state_dict().items()}
    parts = [sd["state_linear.weight"], sd["state_linear.bias"], sd["action_linear.weight"], sd["action_linear.bias"],
             sd["timestep_embed.weight"], sd["stacked_layer_norm.weight"], sd["stacked_layer_norm.bias"]]
    for l in range(desc.num_layers):
        p = f"gpt.layers.{l}."
        parts += [sd[p + "layer_norm_0.weight"], sd[p + "layer_norm_0.bias"], sd[p + "attention.in_linear.weight"],
                  sd[p + "attention.in_linear.bias"], sd[p + "attention.out_linear.weight"],
                  sd[p + "attention.out_linear.bias"], sd[p + "layer_norm_1.weight"], sd[p + "layer_norm_1.bias"],
                  sd[p + "mlp.in_linear.weight"], sd[p + "mlp.in_linear.bias"], sd[p + "mlp.out_linear.weight"],
                  sd[p + "mlp.out_linear.bias"]]
    parts += [sd["gpt.layer_norm.weight"], sd["gpt.layer_norm.bias"], sd["pref_linear.weight"][-1],
              sd["pref_linear.bias"][-1:].clone(), sd["pref_linear.weight"], sd["pref_linear.bias"]]
    flat = torch.cat([p.reshape(-1) for p in parts])
    n = C.c_int64()
    _lib.check(_lib.load().iqlb_pt_param_count(C.byref(desc), C.byref(n)), "iqlb_pt_param_count")
    if flat.numel() != n.value:
        raise ValueError(f"PT weight packing mismatch: {flat.numel()} floats, library expects {n.value}")
    pad = (-flat.numel()) % 4
    if pad:
        flat = torch.cat([flat, torch.zeros(pad)])
    return flat
