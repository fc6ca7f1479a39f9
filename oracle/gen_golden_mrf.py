"""Generate tests/golden/mrf_train.npz by running the UNMODIFIED reference MRTrainerF with an LCFModel prior
(authoring container only).

TEST INFRASTRUCTURE ONLY.  `python -m oracle.gen_golden_mrf` imports `optbnn/training/training.py` (MRTrainerF),
`optbnn/gp/models/model.py` (LCFModel) and `optbnn/bnn/nets/mlp.py` (MLP) from /root/reference, trains K steps on seeded
synthetic preference batches and measurement sets (CPU, fp32 net, fp64 prior algebra as the reference does) and stores
inputs, initial weights, first-step gradients (likelihood + functional prior), per-step logged energy / accuracy, the
prior's log-density at the updated weights, an evaluation() result and the final weights and Adam moments.
Cases: a ReLU net with a diagonal weight covariance, a tanh net with a full covariance, a non-zero prior mean and
bt_pool="sum".
"""
import os
import sys

import numpy as np
import torch

from . import ref_import
from .mr_train_oracle import synth_batch, synth_features

CASES = {
    # name: (S, A, widths, activation, B, T, K, bt_pool, n_meas, n_concepts, full_cov, seed)
    "relu_diag": (9, 3, [64, 64], "relu", 12, 8, 4, "mean", 40, 6, False, 21),
    "tanh_full": (5, 3, [48, 32], "tanh", 7, 6, 3, "sum", 25, 5, True, 22),
}


def prior_arrays(rng, d, full):
    if full:
        a = rng.standard_normal((d, d))
        sigma = a @ a.T / d + 0.5 * np.eye(d)
        p_mean = rng.standard_normal(d) * 0.3
    else:
        sigma = np.diag(rng.uniform(0.5, 2.0, d))
        p_mean = np.zeros(d)
    return p_mean.astype(np.float32), sigma.astype(np.float32)  # LCFModel keeps float32 tensors (model.py:231-243)


def main(out_path):
    ref_import.load_reference_iql()
    from optbnn.bnn.nets.mlp import MLP
    from optbnn.gp.models.model import LCFModel
    from optbnn.training.training import MRTrainerF

    out = {}
    for name, (S, A, widths, act, B, T, K, pool, n_meas, d, full, seed) in CASES.items():
        torch.manual_seed(seed)
        net = MLP(S + A, 1, widths, act)
        with torch.no_grad():
            for lin in list(net.layers) + [net.output_layer]:
                lin.b.normal_(0, 0.1)
        rng = np.random.default_rng(seed)
        p_mean, sigma = prior_arrays(rng, d, full)
        fv = lambda X, device, d=d: torch.from_numpy(synth_features(X.detach().cpu().numpy(), d))  # noqa: E731
        prior = LCFModel(sigma.copy(), fv, "cpu", p_mean=p_mean.copy())
        opt = torch.optim.Adam(net.parameters(), lr=3e-4)
        N = B * K
        tr = MRTrainerF(net, opt, num_datapoints=N, prior=prior, device="cpu", bt_pool=pool)
        batches = [synth_batch(rng, B, T, S, A, ragged=True) for _ in range(K + 1)]
        meas = [rng.standard_normal((n_meas, S + A)).astype(np.float32) for _ in range(K)]
        w0 = [p.detach().numpy().copy() for p in net.parameters()]
        losses, accs, priors = [], [], []
        for k in range(K):
            log = tr.train([torch.from_numpy(x) for x in batches[k]], meas[k])
            if k == 0:
                g0 = [p.grad.detach().numpy().copy() for p in net.parameters()]
            with torch.no_grad():
                xm = torch.from_numpy(meas[k])
                priors.append(float(prior.log_prior(xm, net(xm))))
            losses.append(log["training_loss"]); accs.append(log["training_acc"])
        ev = tr.evaluation([torch.from_numpy(x) for x in batches[K]])
        params = list(net.parameters())
        out[f"{name}/meta"] = np.array([S, A, B, T, K, n_meas, d, N, len(widths)] + widths, dtype=np.int64)
        out[f"{name}/act"] = np.array(act)
        out[f"{name}/pool"] = np.array(pool)
        out[f"{name}/p_mean"] = p_mean
        out[f"{name}/sigma"] = sigma
        for i, w in enumerate(w0):
            out[f"{name}/w0_{i}"] = w
            out[f"{name}/g0_{i}"] = g0[i]
            out[f"{name}/wK_{i}"] = params[i].detach().numpy().copy()
            out[f"{name}/mK_{i}"] = opt.state[params[i]]["exp_avg"].numpy().copy()
            out[f"{name}/vK_{i}"] = opt.state[params[i]]["exp_avg_sq"].numpy().copy()
        for k, b in enumerate(batches):
            for j in (0, 1, 3, 4, 5, 7, 8):
                out[f"{name}/b{k}_{j}"] = b[j]
        for k, m in enumerate(meas):
            out[f"{name}/meas{k}"] = m
        out[f"{name}/loss"] = np.array(losses, dtype=np.float64)
        out[f"{name}/acc"] = np.array(accs, dtype=np.float64)
        out[f"{name}/log_prior"] = np.array(priors, dtype=np.float64)
        out[f"{name}/eval"] = np.array([ev["eval_loss"], ev["eval_acc"]], dtype=np.float64)
        print(name, losses, accs, priors, ev)
    np.savez_compressed(out_path, **out)
    print("wrote", out_path, os.path.getsize(out_path), "bytes")


def load_case(z, name):
    meta = z[f"{name}/meta"]
    S, A, B, T, K, n_meas, d, N, depth = (int(x) for x in meta[:9])
    widths = [int(x) for x in meta[9:9 + depth]]
    n_par = 2 * (depth + 1)
    ts = np.tile(np.arange(T, dtype=np.int64), (B, 1))
    batches = []
    for k in range(K + 1):
        g = {j: z[f"{name}/b{k}_{j}"] for j in (0, 1, 3, 4, 5, 7, 8)}
        batches.append((g[0], g[1], ts, g[3], g[4], g[5], ts, g[7], g[8]))
    lst = lambda p: [z[f"{name}/{p}_{i}"] for i in range(n_par)]  # noqa: E731
    return dict(S=S, A=A, B=B, T=T, K=K, n_meas=n_meas, d=d, N=N, widths=widths, act=str(z[f"{name}/act"]),
                pool=str(z[f"{name}/pool"]), p_mean=z[f"{name}/p_mean"], sigma=z[f"{name}/sigma"],
                w0=lst("w0"), g0=lst("g0"), wK=lst("wK"), mK=lst("mK"), vK=lst("vK"), batches=batches,
                meas=[z[f"{name}/meas{k}"] for k in range(K)], loss=z[f"{name}/loss"], acc=z[f"{name}/acc"],
                log_prior=z[f"{name}/log_prior"], eval=z[f"{name}/eval"])


if __name__ == "__main__":
    here = os.path.dirname(os.path.abspath(__file__))
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(here, "..", "tests", "golden", "mrf_train.npz"))
