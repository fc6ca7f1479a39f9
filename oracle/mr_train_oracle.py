"""CPU restatement of the MR preference-model training step.

TEST INFRASTRUCTURE ONLY: imported by `tests/` (and `__graft_entry__.smoke()`), never by the product path.

Follows, in numpy fp32 with a hand-written backward pass:
  * `MRTrainer.train` / `.evaluation`  gp_reward-priors/optbnn/training/training.py:27-72, 74-114
  * `bt_pool_logit`                    gp_reward-priors/optbnn/utils/util.py:356-364
  * `RewardMLP` (scaled-variance Linear) optbnn/bnn/nets/mlp.py:73-111, optbnn/bnn/layers/linear.py:34-47
  * `torch.nn.CrossEntropyLoss(reduction="mean")` with probability targets
  * `torch.optim.Adam` (optim/adam.py:347-547), lr only, as built in scripts_mr/run_mr_training.py:105
Pinned to the unmodified reference by `tests/golden/mr_train.npz` (`oracle/gen_golden_mr.py`).
"""
import math
from typing import Dict, List, Sequence, Tuple

import numpy as np

from .relabel_oracle import _act

f32 = np.float32


def _act_grad(name: str, z: np.ndarray) -> np.ndarray:
    if name == "relu":
        return (z > 0).astype(f32)
    if name == "tanh":
        return (f32(1) - np.tanh(z) ** 2).astype(f32)
    if name == "cos":
        return -np.sin(z)
    if name == "sin":
        return np.cos(z)
    if name == "linear":
        return np.ones_like(z)
    if name == "rbf":
        return f32(-2) * z * np.exp(-z * z)
    if name == "swish":
        s = f32(1) / (f32(1) + np.exp(-z))
        return s * (f32(1) + z * (f32(1) - s))
    if name == "softplus":
        return np.where(z > 20, f32(1), f32(1) / (f32(1) + np.exp(-np.minimum(z, f32(20))))).astype(f32)
    if name == "leaky_relu":
        return np.where(z > 0, f32(1), f32(0.01)).astype(f32)
    raise ValueError(name)


def mr_forward(weights: Sequence[np.ndarray], batch, activation: str, bt_pool: str = "mean"):
    """Returns (logits [B,2], cache) for batch = (s1, a1, ts1, m1, s2, a2, ts2, m2, labels) (training.py:28-60)."""
    s1, a1, _, m1, s2, a2, _, m2, labels = batch
    B, T, _ = s1.shape
    X = np.concatenate([np.concatenate([s1, a1], -1).reshape(B * T, -1),
                        np.concatenate([s2, a2], -1).reshape(B * T, -1)], 0).astype(f32)
    n_layers = len(weights) // 2
    hs, zs = [X], []
    h = X
    for l in range(n_layers):
        w, b = weights[2 * l].astype(f32), weights[2 * l + 1].astype(f32)
        z = (h @ w) / f32(math.sqrt(w.shape[0])) + b  # linear.py:43-47
        if l < n_layers - 1:
            zs.append(z)
            h = _act(activation, z).astype(f32)
            hs.append(h)
    pred = z.reshape(2, B, T)
    masks = np.stack([m1, m2]).astype(f32)
    pm = pred * masks
    s = np.nansum(pm, axis=2)
    if bt_pool == "mean":
        n = np.maximum(np.nansum(masks, axis=2), f32(1))
    elif bt_pool == "sum":
        n = np.ones_like(s)
    else:
        raise ValueError(f"bt_pool_logit: mode must be 'sum' or 'mean', got {bt_pool!r}")
    logits = (s / n).T.astype(f32)  # [B, 2]
    return logits, dict(hs=hs, zs=zs, masks=masks, n=n, nan=np.isnan(pm), B=B, T=T)


def ce_loss_acc(logits: np.ndarray, labels: np.ndarray) -> Tuple[float, float, np.ndarray]:
    mx = logits.max(1, keepdims=True)
    e = np.exp(logits - mx)
    lse = mx + np.log(e.sum(1, keepdims=True))
    logp = logits - lse
    loss = float(np.mean(-(labels * logp).sum(1)))
    acc = float(np.mean(np.argmax(logits, 1) == np.argmax(labels, 1)))
    dlogits = ((e / e.sum(1, keepdims=True)) * labels.sum(1, keepdims=True) - labels) / f32(logits.shape[0])
    return loss, acc, dlogits.astype(f32)


def mr_loss_and_grads(weights, batch, activation: str, bt_pool: str = "mean"):
    labels = batch[8].astype(f32)
    logits, c = mr_forward(weights, batch, activation, bt_pool)
    loss, acc, dlogits = ce_loss_acc(logits, labels)
    B, T = c["B"], c["T"]
    dpred = dlogits.T[:, :, None] * c["masks"] / c["n"][:, :, None]  # [2, B, T]
    dpred = np.where(c["nan"], f32(0), dpred).astype(f32).reshape(2 * B * T, 1)
    n_layers = len(weights) // 2
    grads: List[np.ndarray] = [None] * len(weights)
    dz = dpred
    for l in range(n_layers - 1, -1, -1):
        w = weights[2 * l].astype(f32)
        sc = f32(1.0 / math.sqrt(w.shape[0]))
        grads[2 * l] = (c["hs"][l].T @ dz) * sc
        grads[2 * l + 1] = dz.sum(0)
        if l > 0:
            dh = (dz @ w.T) * sc
            dz = (dh * _act_grad(activation, c["zs"][l - 1])).astype(f32)
    return loss, acc, grads


def adam_update(weights, grads, m, v, step: int, lr=3e-4, beta1=0.9, beta2=0.999, eps=1e-8):
    """In place; `step` is the count after this update (adam.py:457,476,529-547)."""
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    step_size = lr / bc1
    bc2_sqrt = math.sqrt(bc2)
    for i, g in enumerate(grads):
        g = g.astype(f32)
        m[i] += (g - m[i]) * f32(1.0 - beta1)
        v[i] *= f32(beta2)
        v[i] += f32(1.0 - beta2) * g * g
        denom = np.sqrt(v[i]) / f32(bc2_sqrt) + f32(eps)
        weights[i] -= f32(step_size) * (m[i] / denom)


class MRTrainerOracle:
    def __init__(self, weights: Sequence[np.ndarray], activation: str, lr: float = 3e-4, bt_pool: str = "mean"):
        self.weights = [np.array(w, dtype=f32) for w in weights]
        self.m = [np.zeros_like(w) for w in self.weights]
        self.v = [np.zeros_like(w) for w in self.weights]
        self.activation, self.lr, self.bt_pool, self.step = activation, lr, bt_pool, 0

    def train(self, batch) -> Dict[str, float]:
        loss, acc, grads = mr_loss_and_grads(self.weights, batch, self.activation, self.bt_pool)
        self.step += 1
        adam_update(self.weights, grads, self.m, self.v, self.step, lr=self.lr)
        return {"training_loss": loss, "training_acc": acc}

    def evaluation(self, batch) -> Dict[str, float]:
        logits, _ = mr_forward(self.weights, batch, self.activation, self.bt_pool)
        loss, acc, _ = ce_loss_acc(logits, batch[8].astype(f32))
        return {"eval_loss": loss, "eval_acc": acc}


def synth_batch(rng: np.random.Generator, B: int, T: int, S: int, A: int, ragged: bool = True):
    """Synthetic preference batch shaped like Pref_H5Dataset items (util.py:139-183): left-aligned 0/1 masks."""
    def seg():
        s = rng.standard_normal((B, T, S)).astype(f32)
        a = rng.uniform(-1, 1, (B, T, A)).astype(f32)
        ln = rng.integers(1, T + 1, B) if ragged else np.full(B, T)
        m = (np.arange(T)[None, :] < ln[:, None]).astype(f32)
        ts = np.tile(np.arange(T, dtype=np.int64), (B, 1))
        return s, a, ts, m
    s1, a1, t1, m1 = seg()
    s2, a2, t2, m2 = seg()
    lab = rng.choice(np.array([[1, 0], [0, 1], [0.5, 0.5]], dtype=f32), B)
    return (s1, a1, t1, m1, s2, a2, t2, m2, lab.astype(f32))


# ------------------------------------------------------------------------------------------------------------------
# MRTrainerF: the MR step plus the functional GP prior of an LCF model (training.py:132-305, gp/models/model.py:202-513)
# Pinned to the unmodified reference by `tests/golden/mrf_train.npz` (`oracle/gen_golden_mrf.py`).
# ------------------------------------------------------------------------------------------------------------------
f64 = np.float64


def synth_features(X: np.ndarray, n_concepts: int = 6) -> np.ndarray:
    """The feature map Phi: X -> R^d the MRTrainerF fixtures use as `function_vect` (an LCF prior takes any callable,
    model.py:213-217): an intercept, the first inputs, and two bounded non-linear columns.  float64 [n, d]."""
    X = np.asarray(X, dtype=f64)
    cols = [np.ones(X.shape[0]), X[:, 0], X[:, 1], np.tanh(X[:, 2]), np.sin(X[:, 0] * X[:, 1]), X[:, -1] ** 2]
    return np.stack(cols[:n_concepts], axis=1)


def woodbury_solve(phi: np.ndarray, residual: np.ndarray, sigma: np.ndarray, jitter: float = 1e-6):
    """alpha = (eps I + Phi Sigma Phi^T)^-1 residual and log|eps I + Phi Sigma Phi^T| through the d x d system
    (model.py:262-325): eps = jitter * max(trace(Sigma Phi^T Phi) / n, 1), A = Sigma^-1 + Phi^T Phi / eps."""
    phi, residual, sigma = np.asarray(phi, f64), np.asarray(residual, f64).reshape(-1), np.asarray(sigma, f64)
    n = phi.shape[0]
    if n == 0:
        return np.zeros(0, f64), 0.0
    ptp = phi.T @ phi
    eps = jitter * max(float(np.einsum("ij,ij->", sigma, ptp)) / n, 1.0)
    A = np.linalg.inv(sigma) + ptp / eps
    LA = np.linalg.cholesky(A)
    b = phi.T @ residual
    z = np.linalg.solve(LA.T, np.linalg.solve(LA, b))
    alpha = (residual - (phi @ z) / eps) / eps
    logdet = 2.0 * np.log(np.diag(LA)).sum() + np.linalg.slogdet(sigma)[1] + n * math.log(eps)
    return alpha, float(logdet)


def lcf_log_prior(phi, f, p_mean, sigma, jitter: float = 1e-6) -> float:
    """log p_GP(f(X_M)) (model.py:422-449)."""
    phi = np.asarray(phi, f64)
    residual = np.asarray(f, f64).reshape(-1) - phi @ np.asarray(p_mean, f64)
    alpha, logdet = woodbury_solve(phi, residual, sigma, jitter)
    n = residual.shape[0]
    return float(-0.5 * (residual @ alpha) - 0.5 * logdet - 0.5 * n * math.log(2.0 * math.pi))


def mlp_rows(weights, X: np.ndarray, activation: str):
    """RewardMLP forward over plain rows -> (f [n], cache) (mlp.py:73-111)."""
    n_layers = len(weights) // 2
    h = np.asarray(X, f32)
    hs, zs = [h], []
    for l in range(n_layers):
        w, b = weights[2 * l].astype(f32), weights[2 * l + 1].astype(f32)
        z = (h @ w) / f32(math.sqrt(w.shape[0])) + b
        if l < n_layers - 1:
            zs.append(z)
            h = _act(activation, z).astype(f32)
            hs.append(h)
    return z.reshape(-1), dict(hs=hs, zs=zs)


def mlp_rows_vjp(weights, cache, cot: np.ndarray, activation: str) -> List[np.ndarray]:
    """J_w^T cot for the rows of `cache` (what torch.autograd.grad(f, params, cot) returns, model.py:496-503)."""
    n_layers = len(weights) // 2
    grads: List[np.ndarray] = [None] * len(weights)
    dz = np.asarray(cot, f32).reshape(-1, 1)
    for l in range(n_layers - 1, -1, -1):
        w = weights[2 * l].astype(f32)
        sc = f32(1.0 / math.sqrt(w.shape[0]))
        grads[2 * l] = (cache["hs"][l].T @ dz) * sc
        grads[2 * l + 1] = dz.sum(0)
        if l > 0:
            dz = (((dz @ w.T) * sc) * _act_grad(activation, cache["zs"][l - 1])).astype(f32)
    return grads


class MRTrainerFOracle(MRTrainerOracle):
    """MRTrainerF.train (training.py:166-262): likelihood gradient, plus J^T alpha / N from the functional prior, Adam, and
    the logged energy U / N = CE - log p_GP(f_new(X_M)) / N evaluated with the updated weights."""

    def __init__(self, weights, activation, num_datapoints, p_mean, sigma, lr=3e-4, bt_pool="mean"):
        super().__init__(weights, activation, lr=lr, bt_pool=bt_pool)
        self.N, self.p_mean, self.sigma = float(num_datapoints), np.asarray(p_mean, f64), np.asarray(sigma, f64)

    def grads(self, batch, x_meas, phi, jitter=1e-6):
        loss, acc, g = mr_loss_and_grads(self.weights, batch, self.activation, self.bt_pool)
        f, cache = mlp_rows(self.weights, x_meas, self.activation)
        alpha, _ = woodbury_solve(phi, f.astype(f64) - np.asarray(phi, f64) @ self.p_mean, self.sigma, jitter)
        vjp = mlp_rows_vjp(self.weights, cache, alpha.astype(f32), self.activation)  # grad_out = alpha.to(f32), :484
        # functional_prior_grad returns -J^T alpha; param.grad.add_(-fg / N) (training.py:236-238)
        g = [(gi.reshape(vi.shape) + vi / f32(self.N)).astype(f32) for gi, vi in zip(g, vjp)]
        return loss, acc, g

    def train(self, batch, x_meas, phi, jitter=1e-6) -> Dict[str, float]:
        loss, acc, g = self.grads(batch, x_meas, phi, jitter)
        self.step += 1
        adam_update(self.weights, g, self.m, self.v, self.step, lr=self.lr)
        f_new, _ = mlp_rows(self.weights, x_meas, self.activation)
        prior = lcf_log_prior(phi, f_new, self.p_mean, self.sigma, jitter)
        return {"training_loss": loss - prior / self.N, "training_acc": acc, "ce_loss": loss, "log_prior": prior}
