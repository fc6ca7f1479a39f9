"""`PTTrainer` of /root/reference/gp_reward-priors/optbnn/training/training.py:308-416 behind the same surface
(`PTTrainer(net, opt, device, bt_pool)`, `.train(batch) -> {"training_loss", "training_acc"}`,
`.evaluation(batch) -> {"eval_loss", "eval_acc"}`, `.state_dict()`, `.load_state_dict()`), SURVEY.md §8f-1.

The whole step - one forward of the preference transformer over both segments in training mode
(optbnn/bnn/nets/pref_trans.py:269-325), `bt_pool_logit` on `weighted_sum` (optbnn/utils/util.py:356-364), soft-label
cross entropy, the backward pass and `torch.optim.Adam` - is `iqlb_pt_train_step` in libiqlb.so (`csrc/pt_train.cu`).
The network's parameters become views of one flat fp32 device buffer in `net.parameters()` order, so `net.state_dict()`,
`best_model.pt` (what `load_pt_reward_model`, iql.py:1312-1336, reads) and `qlearning_dataset_pt` see the trained weights
without a copy; Adam's moments live in two more flat buffers and are exposed through `opt.state` in the reference's
format.  `net` is this package's `reward_models.pref_transformer.PT` or the reference's `optbnn` `PT` (same parameter
names).  Dropout: the reference draws its masks from torch's generator, which cannot be mirrored - the step applies the
Philox masks documented in include/iqlb.h (`seed` below).  No CPU path.
"""
import ctypes as C
from typing import Any, Dict, List

import torch
import torch.nn as nn

from . import _lib
from .mr_training import fit_mr
from .reward_models.pref_transformer import pt_desc_from_model

TensorBatch = List[torch.Tensor]


def pt_parameter_names(num_layers: int) -> List[str]:
    """`[n for n, _ in PT(...).named_parameters()]` (pref_trans.py:246-267, 197-214, 157-172, 76-79, 104-108)."""
    names = ["state_linear.weight", "state_linear.bias", "action_linear.weight", "action_linear.bias", "timestep_embed.weight",
             "stacked_layer_norm.weight", "stacked_layer_norm.bias"]
    for l in range(num_layers):
        p = f"gpt.layers.{l}."
        names += [p + "layer_norm_0.weight", p + "layer_norm_0.bias", p + "attention.in_linear.weight",
                  p + "attention.in_linear.bias", p + "attention.out_linear.weight", p + "attention.out_linear.bias",
                  p + "layer_norm_1.weight", p + "layer_norm_1.bias", p + "mlp.in_linear.weight", p + "mlp.in_linear.bias",
                  p + "mlp.out_linear.weight", p + "mlp.out_linear.bias"]
    return names + ["gpt.layer_norm.weight", "gpt.layer_norm.bias", "pref_linear.weight", "pref_linear.bias"]


def _dropout_ps(net):
    """(embd, attention, after attention.out_linear, after mlp.out_linear).  The reference's modules carry them
    (`GPT2Model.dropout`, `GPT2SelfAttention.attn_dropout / .resid_dropout`, `GPT2MLP.resid_dropout`); this package's
    parameter container keeps the constructor's values (`reward_models/pref_transformer.py`)."""
    gpt = net.gpt
    if hasattr(gpt, "dropout"):
        lay = gpt.layers[0]
        ps = (gpt.dropout.p, lay.attention.attn_dropout.p, lay.attention.resid_dropout.p, lay.mlp.resid_dropout.p)
        for other in gpt.layers[1:]:
            if (other.attention.attn_dropout.p, other.attention.resid_dropout.p, other.mlp.resid_dropout.p) != ps[1:]:
                raise NotImplementedError("PTTrainer: the blocks must share their dropout probabilities")
        return tuple(float(p) for p in ps)
    return (float(getattr(net, "embd_dropout_p", 0.0)), float(getattr(net, "attn_dropout_p", 0.0)),
            float(getattr(net, "resid_dropout_p", 0.0)), float(getattr(net, "resid_dropout_p", 0.0)))


class PTTrainer:
    def __init__(self, net: nn.Module, opt: torch.optim.Optimizer, device: str = "cuda", bt_pool: str = "mean",
                 max_pairs: int = 0, precision: str = "fp32", seed: int = None):
        """`precision`: "fp32" (FFMA GEMMs, the parity mode) or "split_bf16" (every Linear of the step - forward, data and
        weight gradients - on the tensor cores, x_hi*w_hi + x_lo*w_hi + x_hi*w_lo with fp32 accumulation).
        `seed`: key of the dropout masks (default `torch.initial_seed()`)."""
        if precision not in ("fp32", "split_bf16"):
            raise ValueError(f"precision must be 'fp32' or 'split_bf16', got {precision!r}")
        dev = torch.device(device)
        if dev.type != "cuda":
            raise RuntimeError("iqlpref_b200 has no CPU path: PTTrainer runs on a CUDA device")
        if bt_pool not in ("mean", "sum"):
            raise ValueError(f"bt_pool_logit: mode must be 'sum' or 'mean', got {bt_pool!r}")
        if not isinstance(opt, torch.optim.Adam) or isinstance(opt, torch.optim.AdamW):
            raise NotImplementedError("PTTrainer: torch.optim.Adam only (scripts_pt/run_pt_training.py:137)")
        g = opt.param_groups[0]
        if len(opt.param_groups) != 1 or g.get("weight_decay", 0) != 0 or g.get("amsgrad", False) or g.get("maximize", False):
            raise NotImplementedError("PTTrainer: one parameter group, no weight decay / amsgrad / maximize")
        self.net, self.opt, self.device, self.bt_pool, self.precision = net, opt, dev, bt_pool, precision
        core = getattr(net, "_orig_mod", net)
        self._core = core
        named = list(core.named_parameters())
        layers = len(core.gpt.layers)
        if [n for n, _ in named] != pt_parameter_names(layers):
            raise ValueError("PTTrainer: the network's parameters are not those of optbnn's PT (pref_trans.py:246-267)")
        params = [p for _, p in named]
        if [id(p) for p in params] != [id(p) for p in g["params"]]:
            raise ValueError("PTTrainer: the optimizer must hold net.parameters() in order")
        self.flat = torch.cat([p.detach().to(dev, torch.float32).reshape(-1) for p in params]).contiguous()
        self.exp_avg = torch.zeros_like(self.flat)
        self.exp_avg_sq = torch.zeros_like(self.flat)
        self._slices = []
        off = 0
        for p in params:
            n = p.numel()
            self._slices.append((off, n, tuple(p.shape)))
            p.data = self.flat[off:off + n].view(p.shape)
            off += n
        for name, buf in core.named_buffers():  # causal_bias: keep the module on one device
            buf.data = buf.data.to(dev)
        self._params = params
        self.step_count = 0
        self.seed = int(torch.initial_seed() if seed is None else seed) & ((1 << 64) - 1)
        self._handle, self._desc = None, None
        self._max_pairs = int(max_pairs)
        self._out2 = torch.zeros(2, dtype=torch.float32, device=dev)
        self._adopt_optimizer_state()

    # ---------------------------------------------------------------- native handle
    def _ensure(self, pairs: int, query_len: int):
        lib = _lib.load()
        if self._handle is not None and pairs <= self._max_pairs and self._desc.query_len == query_len:
            return lib
        self._release()
        self._max_pairs = max(pairs, self._max_pairs)
        desc = pt_desc_from_model(self._core, query_len)
        n = C.c_int64()
        _lib.check(lib.iqlb_pt_train_param_count(C.byref(desc), C.byref(n)), "iqlb_pt_train_param_count")
        if n.value != self.flat.numel():
            raise RuntimeError(f"parameter count {self.flat.numel()} != {n.value} expected by the library")
        h = C.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(lib.iqlb_pt_trainer_create(C.byref(desc), 1 if self.bt_pool == "mean" else 0, self._max_pairs,
                                                  C.byref(h)), "iqlb_pt_trainer_create")
        _lib.check(lib.iqlb_pt_trainer_set_precision(h, 1 if self.precision == "split_bf16" else 0),
                   "iqlb_pt_trainer_set_precision")
        self._handle, self._desc = h, desc
        return lib

    def _release(self):
        if self._handle is not None:
            _lib.load().iqlb_pt_trainer_destroy(self._handle)
            self._handle = None

    def __del__(self):
        try:
            self._release()
        except Exception:
            pass

    def _unpack(self, batch: TensorBatch):
        s1, a1, t1, m1, s2, a2, t2, m2, labels = batch
        if s1.dim() != 3 or a1.dim() != 3:
            raise ValueError("states / actions must be [B, T, dim]")
        B, T, S = s1.shape
        A = a1.shape[2]
        f = lambda t: t.to(self.device, torch.float32).contiguous()
        i = lambda t: t.to(self.device, torch.int32).contiguous()  # training.py:343 `.to(torch.int)`
        ts = [f(s1), f(a1), i(t1), f(m1), f(s2), f(a2), i(t2), f(m2), f(labels)]
        if ts[4].shape != (B, T, S) or ts[5].shape != (B, T, A) or ts[8].shape != (B, 2) or \
                any(t.shape != (B, T) for t in (ts[2], ts[3], ts[6], ts[7])):
            raise ValueError("inconsistent preference batch shapes")
        if S != self._core.state_linear.weight.shape[1] or A != self._core.action_linear.weight.shape[1]:
            raise RuntimeError(f"mat1 and mat2 shapes cannot be multiplied (state_dim {S} / action_dim {A} vs the network's "
                               f"{self._core.state_linear.weight.shape[1]} / {self._core.action_linear.weight.shape[1]})")
        n_ts = self._core.timestep_embed.weight.shape[0]
        lo = min(int(ts[2].min()), int(ts[6].min()))
        hi = max(int(ts[2].max()), int(ts[6].max()))
        if lo < 0 or hi >= n_ts:
            raise IndexError("index out of range in self")  # nn.Embedding's error (pref_trans.py:276)
        return ts, int(B), int(T)

    def _invalidate_packed(self):
        # the relabel path caches a packed copy of the weights keyed on tensor versions; the step writes through raw pointers
        if getattr(self._core, "_iqlb_packed", None) is not None:
            object.__setattr__(self._core, "_iqlb_packed", None)

    # ---------------------------------------------------------------- training.py:322-368
    def train(self, batch: TensorBatch, return_grad: bool = False, sync: bool = True) -> Dict[str, float]:
        ts, B, T = self._unpack(batch)
        if not self.net.training:
            self.net.train()
        lib = self._ensure(B, T)
        g = self.opt.param_groups[0]
        ps = _dropout_ps(self._core)
        _lib.check(lib.iqlb_pt_trainer_set_dropout(self._handle, *ps, C.c_uint64(self.seed)), "iqlb_pt_trainer_set_dropout")
        grad = torch.empty_like(self.flat) if return_grad else None
        self.step_count += 1
        with torch.cuda.device(self.device):
            _lib.check(lib.iqlb_pt_train_step(self._handle, _lib.ptr(self.flat), _lib.ptr(self.exp_avg),
                                              _lib.ptr(self.exp_avg_sq), *[_lib.ptr(t) for t in ts], B, float(g["lr"]),
                                              float(g["betas"][0]), float(g["betas"][1]), float(g["eps"]), self.step_count, 1,
                                              _lib.ptr(self._out2), _lib.ptr(grad), _lib.stream_ptr(self.device)),
                       "iqlb_pt_train_step")
        self._invalidate_packed()
        if sync:
            loss, acc = self._out2.tolist()  # the reference's loss.item() sync (training.py:361-366)
        else:
            snap = self._out2.clone()
            loss, acc = snap[0], snap[1]
        log = {"training_loss": loss, "training_acc": acc}
        if return_grad:
            log["grad"] = grad
        return log

    def loss_and_grad(self, batch: TensorBatch, training: bool = False):
        """Loss, accuracy and the flat gradient without an optimizer step (diagnostics / tests); `training` applies the
        dropout masks of Adam step `step_count + 1`."""
        ts, B, T = self._unpack(batch)
        lib = self._ensure(B, T)
        ps = _dropout_ps(self._core)
        _lib.check(lib.iqlb_pt_trainer_set_dropout(self._handle, *ps, C.c_uint64(self.seed)), "iqlb_pt_trainer_set_dropout")
        grad = torch.empty_like(self.flat)
        with torch.cuda.device(self.device):
            _lib.check(lib.iqlb_pt_train_step(self._handle, _lib.ptr(self.flat), None, None, *[_lib.ptr(t) for t in ts], B, 0.0,
                                              0.9, 0.999, 1e-8, self.step_count + 1, 1 if training else 0,
                                              _lib.ptr(self._out2), _lib.ptr(grad), _lib.stream_ptr(self.device)),
                       "iqlb_pt_train_step")
        loss, acc = self._out2.tolist()
        return loss, acc, grad

    # ---------------------------------------------------------------- training.py:370-405
    def evaluation(self, batch: TensorBatch) -> Dict[str, float]:
        ts, B, T = self._unpack(batch)
        if self.net.training:
            self.net.eval()
        lib = self._ensure(B, T)
        with torch.cuda.device(self.device):
            _lib.check(lib.iqlb_pt_eval(self._handle, _lib.ptr(self.flat), *[_lib.ptr(t) for t in ts], B,
                                        _lib.ptr(self._out2), None, _lib.stream_ptr(self.device)), "iqlb_pt_eval")
        loss, acc = self._out2.tolist()
        return {"eval_loss": loss, "eval_acc": acc}

    def grad_dict(self, flat_grad: torch.Tensor) -> Dict[str, torch.Tensor]:
        names = pt_parameter_names(len(self._core.gpt.layers))
        return {n: flat_grad[off:off + k].view(shape) for n, (off, k, shape) in zip(names, self._slices)}

    # ---------------------------------------------------------------- optimizer state in the reference's format
    def _publish_optimizer_state(self):
        for p, (off, n, shape) in zip(self._params, self._slices):
            self.opt.state[p] = {"step": torch.tensor(float(self.step_count)),
                                 "exp_avg": self.exp_avg[off:off + n].view(shape),
                                 "exp_avg_sq": self.exp_avg_sq[off:off + n].view(shape)}

    def _adopt_optimizer_state(self):
        steps = set()
        for p, (off, n, shape) in zip(self._params, self._slices):
            st = self.opt.state.get(p)
            if not st:
                continue
            self.exp_avg[off:off + n].copy_(st["exp_avg"].to(self.device, torch.float32).reshape(-1))
            self.exp_avg_sq[off:off + n].copy_(st["exp_avg_sq"].to(self.device, torch.float32).reshape(-1))
            steps.add(int(float(st["step"])))
        if len(steps) > 1:
            raise ValueError("PTTrainer: parameters with different Adam step counts")
        if steps:
            self.step_count = steps.pop()

    def state_dict(self) -> Dict[str, Any]:  # training.py:407-411
        self._publish_optimizer_state()
        return {"net": self.net.state_dict(), "optimizer": self.opt.state_dict()}

    def load_state_dict(self, state_dict: Dict[str, Any]):  # training.py:413-415
        self.net.load_state_dict(state_dict["net"])  # copies into the views of the flat buffer
        self.opt.load_state_dict(state_dict["optimizer"])
        self._adopt_optimizer_state()
        self._invalidate_packed()


def fit_pt(model: PTTrainer, training_batches, val_batches, epochs: int, eval_every: int = 1, checkpoints_path: str = None,
           criteria_key: str = "acc", log=None):
    """The epoch loop of `scripts_pt/run_pt_training.py:146-239` around `PTTrainer`: epoch 0 evaluates only, every
    `eval_every` epochs the validation pass runs and `best_model.pt` follows the reference's criteria (the same rules as
    the MR script; the PT script writes no per-epoch snapshots)."""
    return fit_mr(model, training_batches, val_batches, epochs, eval_every, checkpoints_path, criteria_key, None, log,
                  snapshots=False)
