"""`MRTrainer` (and `MRTrainerF`, :132-305, below) of /root/reference/gp_reward-priors/optbnn/training/training.py:11-129 behind the same surface
(`MRTrainer(net, opt, num_datapoints, device, bt_pool)`, `.train(batch) -> {"training_loss", "training_acc"}`,
`.evaluation(batch) -> {"eval_loss", "eval_acc"}`, `.state_dict()`, `.load_state_dict()`), SURVEY.md §8f-1.

The whole step - forward over both segments, `bt_pool_logit` (optbnn/utils/util.py:356-364), soft-label cross
entropy, backward, `torch.optim.Adam` - is `iqlb_mr_train_step` in libiqlb.so (`csrc/mr_train.cu`).  The network's
parameters become views of one flat fp32 device buffer in the member layout the relabel kernels read
(`W0 b0 ... Wout bout`, W stored (n_in, n_out)), so a trained net goes to `ensemble_predict` / `qlearning_dataset_mr`
without a copy and `net.state_dict()` / `checkpoint_{epoch}.pt` files keep the reference's keys.  Adam's moments live
in two more flat buffers and are exposed through `opt.state` as views in the reference's format.  No CPU path.
"""
import ctypes as C
from typing import Any, Dict, List, Sequence

import torch
import torch.nn as nn

from . import _lib

TensorBatch = List[torch.Tensor]


def _linears(net) -> List[nn.Module]:
    net = getattr(net, "_orig_mod", net)
    return list(net.layers) + [net.output_layer]


class MRTrainer:
    def __init__(self, net: nn.Module, opt: torch.optim.Optimizer, num_datapoints: int, device: str = "cuda",
                 bt_pool: str = "mean", max_rows: int = 0, precision: str = "fp32"):
        """`precision`: "fp32" (FFMA GEMMs, the parity mode) or "split_bf16" (the step's GEMMs on the tensor cores,
        every product as x_hi*w_hi + x_lo*w_hi + x_hi*w_lo with fp32 accumulation, ~1e-5 relative)."""
        if precision not in ("fp32", "split_bf16"):
            raise ValueError(f"precision must be 'fp32' or 'split_bf16', got {precision!r}")
        self.precision = precision
        dev = torch.device(device)
        if dev.type != "cuda":
            raise RuntimeError("iqlpref_b200 has no CPU path: MRTrainer runs on a CUDA device")
        if bt_pool not in ("mean", "sum"):
            raise ValueError(f"bt_pool_logit: mode must be 'sum' or 'mean', got {bt_pool!r}")
        if not isinstance(opt, torch.optim.Adam) or isinstance(opt, torch.optim.AdamW):
            raise NotImplementedError("MRTrainer: torch.optim.Adam only (scripts_mr/run_mr_training.py:105)")
        g = opt.param_groups[0]
        if len(opt.param_groups) != 1 or g.get("weight_decay", 0) != 0 or g.get("amsgrad", False) or g.get("maximize", False):
            raise NotImplementedError("MRTrainer: one parameter group, no weight decay / amsgrad / maximize")
        self.net, self.opt, self.device, self.bt_pool, self.num_datapoints = net, opt, dev, bt_pool, num_datapoints
        lins = _linears(net)
        from .relabel import _activation_name
        self.activation = _activation_name(getattr(net, "_orig_mod", net))
        if lins[-1].W.shape[1] != 1:
            raise ValueError("reward MLP must have output_dim == 1")
        self.d_in = int(lins[0].W.shape[0])
        self.widths = [int(l.W.shape[1]) for l in lins[:-1]]
        params = [p for l in lins for p in (l.W, l.b)]
        if [id(p) for p in params] != [id(p) for p in g["params"]]:
            raise ValueError("MRTrainer: the optimizer must hold net.parameters() in order")
        # one flat buffer; the module's parameters become views of it
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
        self._params = params
        self.step_count = 0
        self._handle = None
        self._max_rows = int(max_rows)
        self._out2 = torch.zeros(2, dtype=torch.float32, device=dev)
        self._adopt_optimizer_state()

    # ---------------------------------------------------------------- native handle
    def _ensure(self, rows: int):
        lib = _lib.load()
        if self._handle is not None and rows <= self._max_rows:
            return lib
        if self._handle is not None:
            lib.iqlb_mr_destroy(self._handle)
            self._handle = None
        self._max_rows = max(rows, self._max_rows)
        h = C.c_void_p()
        widths = (C.c_int32 * len(self.widths))(*self.widths)
        with torch.cuda.device(self.device):
            _lib.check(lib.iqlb_mr_create(self.d_in, widths, len(self.widths), _lib.ACTIVATIONS[self.activation],
                                          1 if self.bt_pool == "mean" else 0, self._max_rows, C.byref(h)),
                       "iqlb_mr_create")
        _lib.check(lib.iqlb_mr_set_precision(h, 1 if self.precision == "split_bf16" else 0), "iqlb_mr_set_precision")
        n = C.c_int64()
        _lib.check(lib.iqlb_mr_param_count(h, C.byref(n)), "iqlb_mr_param_count")
        assert n.value == self.flat.numel()
        self._handle = h
        return lib

    def __del__(self):
        try:
            if self._handle is not None:
                _lib.load().iqlb_mr_destroy(self._handle)
                self._handle = None
        except Exception:
            pass

    def _unpack(self, batch: TensorBatch):
        s1, a1, _, m1, s2, a2, _, m2, labels = batch
        if s1.dim() != 3 or a1.dim() != 3:
            raise ValueError("states / actions must be [B, T, dim]")
        B, T, S = s1.shape
        A = a1.shape[2]
        prep = lambda t: t.to(self.device, torch.float32).contiguous()
        ts = [prep(t) for t in (s1, a1, m1, s2, a2, m2, labels)]
        if ts[3].shape != (B, T, S) or ts[4].shape != (B, T, A) or ts[2].shape != (B, T) or ts[5].shape != (B, T) \
                or ts[6].shape != (B, 2):
            raise ValueError("inconsistent preference batch shapes")
        if S + A != self.d_in:
            raise RuntimeError(f"mat1 and mat2 shapes cannot be multiplied ({2 * B * T}x{S + A} and {self.d_in}x{self.widths[0]})")
        return ts, int(B), int(T), int(S), int(A)

    # ---------------------------------------------------------------- training.py:27-72
    def train(self, batch: TensorBatch, return_grad: bool = False, sync: bool = True) -> Dict[str, float]:
        """`sync=False` skips the reference's per-step `loss.item()`: the log holds 0-dim CUDA tensors (a snapshot of
        this step's loss / accuracy) and the host runs ahead of the device."""
        ts, B, T, S, A = self._unpack(batch)
        if not self.net.training:
            self.net.train()
        lib = self._ensure(2 * B * T)
        g = self.opt.param_groups[0]
        grad = torch.empty_like(self.flat) if return_grad else None
        self.step_count += 1
        with torch.cuda.device(self.device):
            _lib.check(lib.iqlb_mr_train_step(self._handle, _lib.ptr(self.flat), _lib.ptr(self.exp_avg),
                                              _lib.ptr(self.exp_avg_sq), *[_lib.ptr(t) for t in ts], B, T, S, A,
                                              float(g["lr"]), float(g["betas"][0]), float(g["betas"][1]),
                                              float(g["eps"]), self.step_count, _lib.ptr(self._out2), _lib.ptr(grad),
                                              _lib.stream_ptr(self.device)), "iqlb_mr_train_step")
        if sync:
            loss, acc = self._out2.tolist()  # the reference's loss.item() sync (training.py:65-70)
        else:
            snap = self._out2.clone()
            loss, acc = snap[0], snap[1]
        log = {"training_loss": loss, "training_acc": acc}
        if return_grad:
            log["grad"] = grad
        return log

    # ---------------------------------------------------------------- training.py:74-114
    def evaluation(self, batch: TensorBatch) -> Dict[str, float]:
        ts, B, T, S, A = self._unpack(batch)
        if self.net.training:
            self.net.eval()
        lib = self._ensure(2 * B * T)
        with torch.cuda.device(self.device):
            _lib.check(lib.iqlb_mr_eval(self._handle, _lib.ptr(self.flat), *[_lib.ptr(t) for t in ts], B, T, S, A,
                                        _lib.ptr(self._out2), None, _lib.stream_ptr(self.device)), "iqlb_mr_eval")
        loss, acc = self._out2.tolist()
        return {"eval_loss": loss, "eval_acc": acc}

    def member(self) -> List[torch.Tensor]:
        """[W0, b0, ..., Wout, bout] views for `ensemble_predict` / snapshot ensembles (no copy)."""
        return [p.detach() for p in self._params]

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
            raise ValueError("MRTrainer: parameters with different Adam step counts")
        if steps:
            self.step_count = steps.pop()

    def state_dict(self) -> Dict[str, Any]:  # training.py:116-120
        self._publish_optimizer_state()
        return {"net": self.net.state_dict(), "optimizer": self.opt.state_dict()}

    def load_state_dict(self, state_dict: Dict[str, Any]):  # training.py:122-124
        self.net.load_state_dict(state_dict["net"])  # copies into the views of the flat buffer
        self.opt.load_state_dict(state_dict["optimizer"])
        self._adopt_optimizer_state()


class MRTrainerF(MRTrainer):
    """`MRTrainerF` of training.py:132-305: the MR step with the functional GP prior of an `LCFModel` in place of a
    weight-space prior.  `train(batch, x_meas, aux_meas=None, jitter=1e-6)` descends U(f) / N = CE + I_0(f) / N: the
    measurement points ride through the step's GEMMs as extra rows, their cotangent alpha / N =
    (K + eps I)^-1 (f(X_M) - m(X_M)) / N comes from the d x d Woodbury system in fp64 (`mr_fprior_kernel`), and the logged
    `training_loss` is CE - log p_GP(f_new(X_M)) / N with the updated weights (training.py:242-250) -
    `iqlb_mr_train_step_f` (`csrc/mr_train.cu`).  `prior` is this package's `gp_prior.LCFModel`, or any object with the
    reference LCFModel's `p_mean`, `p_covariance` and `function_vect` (it is re-wrapped).  `evaluation`, `state_dict`,
    `load_state_dict` are MRTrainer's (the reference's bodies are identical, training.py:264-305)."""

    def __init__(self, net, opt, num_datapoints: int, prior, device: str = "cuda", bt_pool: str = "mean",
                 max_rows: int = 0, precision: str = "fp32"):
        super().__init__(net, opt, num_datapoints, device=device, bt_pool=bt_pool, max_rows=max_rows, precision=precision)
        from .gp_prior import LCFModel
        if prior is None:
            raise ValueError("MRTrainerF: the functional prior is required (training.py:141)")
        if not isinstance(prior, LCFModel):
            prior = LCFModel(prior.p_covariance.detach().cpu(), prior.function_vect, self.device,
                             p_mean=prior.p_mean.detach().cpu(), name=getattr(prior, "name", None))
        self.prior = prior
        self._prior_out = torch.zeros(2, dtype=torch.float64, device=self.device)

    def train(self, batch: TensorBatch, x_meas, aux_meas=None, jitter: float = 1e-6, return_grad: bool = False,
              sync: bool = True) -> Dict[str, float]:
        ts, B, T, S, A = self._unpack(batch)
        if not self.net.training:
            self.net.train()
        x_meas_t = torch.as_tensor(x_meas, dtype=torch.float32, device=self.device).contiguous()  # training.py:219
        aux_t = torch.as_tensor(aux_meas, dtype=torch.float32, device=self.device) if aux_meas is not None else None
        if x_meas_t.ndim != 2 or x_meas_t.shape[1] != self.d_in:
            raise RuntimeError(f"mat1 and mat2 shapes cannot be multiplied ({tuple(x_meas_t.shape)} and {self.d_in}x{self.widths[0]})")
        n_meas = int(x_meas_t.shape[0])
        if n_meas == 0:
            raise ValueError("MRTrainerF: empty measurement set")
        phi = self.prior.features64(x_meas_t, aux_t)
        if phi.shape[0] != n_meas:
            raise ValueError("function_vect must return one row per measurement point")
        rows = 2 * B * T + n_meas
        lib = self._ensure(rows + (rows & 1))
        g = self.opt.param_groups[0]
        grad = torch.empty_like(self.flat) if return_grad else None
        self.step_count += 1
        pr = self.prior
        with torch.cuda.device(self.device):
            _lib.check(lib.iqlb_mr_train_step_f(
                self._handle, _lib.ptr(self.flat), _lib.ptr(self.exp_avg), _lib.ptr(self.exp_avg_sq),
                *[_lib.ptr(t) for t in ts], B, T, S, A, float(g["lr"]), float(g["betas"][0]), float(g["betas"][1]),
                float(g["eps"]), self.step_count, _lib.ptr(x_meas_t), _lib.ptr(phi), n_meas, pr.n_concepts,
                _lib.ptr(pr._p_mean64), _lib.ptr(pr._sigma64), _lib.ptr(pr._sigma_inv64), pr._logdet_sigma, float(jitter),
                float(self.num_datapoints), _lib.ptr(self._out2), _lib.ptr(self._prior_out), _lib.ptr(grad),
                _lib.stream_ptr(self.device)), "iqlb_mr_train_step_f")
        if sync:
            loss, acc = self._out2.tolist()
            prior_val = float(self._prior_out[0])
            log = {"training_loss": loss - prior_val / self.num_datapoints, "training_acc": acc}
        else:
            snap, psnap = self._out2.clone(), self._prior_out.clone()
            log = {"training_loss": snap[0].double() - psnap[0] / self.num_datapoints, "training_acc": snap[1]}
        if return_grad:
            log["grad"] = grad
            log["ce_loss"] = float(self._out2[0])
            log["log_prior"] = float(self._prior_out[0])
        return log


def fit_mr(model: MRTrainer, training_batches, val_batches, epochs: int, eval_every: int = 5,
           checkpoints_path: str = None, criteria_key: str = "acc", activations: str = None, log=None,
           snapshots: bool = True):
    """The epoch loop of `scripts_mr/run_mr_training.py:117-215` around `MRTrainer`: epoch 0 evaluates only, every
    `eval_every` epochs the validation pass runs and `checkpoint_{epoch}.pt` is written (the snapshot ensemble that
    `qlearning_dataset_mr_ensemble` reads), `best_model.pt` follows the reference's criteria ("acc": higher accuracy,
    ties broken by lower loss; "loss": the mirror image).  `training_batches` / `val_batches` are re-iterable
    collections of 9-tensor batches (a `DataLoader` over `Pref_H5Dataset`, or a list).  With `activations` given, a
    `config.yaml` holding it is written beside the snapshots (what `_mr_activations`, iql.py:1339-1342, reads).
    Returns the per-epoch metric dicts (what the reference sends to wandb)."""
    import os

    import numpy as np
    if criteria_key not in ("acc", "loss"):
        raise ValueError(f"criteria_key must be 'acc' or 'loss', got {criteria_key!r}")
    if checkpoints_path is not None:
        os.makedirs(checkpoints_path, exist_ok=True)
        if activations is not None:
            import yaml
            with open(os.path.join(checkpoints_path, "config.yaml"), "w") as f:
                yaml.safe_dump({"activations": activations}, f)

    def save(name):
        if checkpoints_path is not None:
            torch.save(model.state_dict(), os.path.join(checkpoints_path, name))

    c_best_epoch, best_acc, best_loss = 0, -np.inf, np.inf
    history = []
    for epoch in range(epochs + 1):
        metrics = {"training_loss": [], "training_acc": [], "best_epoch": c_best_epoch, "eval_loss": [], "eval_acc": [],
                   f"eval_{criteria_key}_best": best_acc if criteria_key == "acc" else best_loss}
        if epoch:
            for batch in training_batches:
                for key, val in model.train(batch).items():
                    metrics[key].append(val)
        else:
            metrics["training_loss"] = np.nan
        if epoch % eval_every == 0:
            for batch in val_batches:
                for key, val in model.evaluation(batch).items():
                    metrics[key].append(val)
            loss, acc = np.mean(metrics["eval_loss"]), np.mean(metrics["eval_acc"])
            if snapshots:
                save(f"checkpoint_{epoch}.pt")
            first, second = (acc, -loss) if criteria_key == "acc" else (-loss, acc)
            best_first, best_second = (best_acc, -best_loss) if criteria_key == "acc" else (-best_loss, best_acc)
            if first > best_first or (first == best_first and second > best_second):
                c_best_epoch = epoch
                metrics["best_epoch"] = c_best_epoch
                save("best_model.pt")
            if first > best_first:  # run_mr_training.py:153-205: each branch updates only what it compared
                best_acc, best_loss = (acc, min(loss, best_loss)) if criteria_key == "acc" else (max(acc, best_acc), loss)
            elif first == best_first:
                if second > best_second:
                    if criteria_key == "acc":
                        best_loss = loss
                    else:
                        best_acc = acc
            else:
                if criteria_key == "acc":
                    best_loss = min(loss, best_loss)
                else:
                    best_acc = max(acc, best_acc)
            metrics[f"eval_{criteria_key}_best"] = best_acc if criteria_key == "acc" else best_loss
        metrics = {k: (float(np.mean(v)) if isinstance(v, list) else v) for k, v in metrics.items()
                   if not (isinstance(v, list) and not len(v))}
        history.append(metrics)
        if log is not None:
            log(metrics, epoch)
    return history
