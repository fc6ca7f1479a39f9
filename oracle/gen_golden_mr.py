"""Generate tests/golden/mr_train.npz by running the UNMODIFIED reference MRTrainer (authoring container only).

TEST INFRASTRUCTURE ONLY.  `python -m oracle.gen_golden_mr` imports
`gp_reward-priors/optbnn/training/training.py` (MRTrainer) and `optbnn/bnn/nets/mlp.py` (MLP) from
/root/reference, trains K steps on seeded synthetic preference batches (CPU, fp32) and stores the inputs,
the initial weights, the first-step gradients, per-step loss / accuracy, an evaluation() result and the final
weights and Adam moments.  Cases: antmaze-shaped ReLU net (37 -> 128x3 -> 1, the eval yaml's shape, short
segments), a tanh net with ragged masks and soft labels, a swish net with bt_pool="sum".
"""
import os
import sys

import numpy as np
import torch

from . import ref_import
from .mr_train_oracle import synth_batch

CASES = {
    # name: (S, A, widths, activation, B, T, K, bt_pool, seed)
    "antmaze_relu": (29, 8, [128, 128, 128], "relu", 16, 12, 4, "mean", 11),
    "ragged_tanh": (5, 3, [64, 32], "tanh", 9, 7, 5, "mean", 12),
    "sum_swish": (4, 2, [48], "swish", 6, 5, 3, "sum", 13),
}


def main(out_path):
    ref_import.load_reference_iql()  # installs the stubs and puts gp_reward-priors on sys.path
    from optbnn.bnn.nets.mlp import MLP
    from optbnn.training.training import MRTrainer

    out = {}
    for name, (S, A, widths, act, B, T, K, pool, seed) in CASES.items():
        torch.manual_seed(seed)
        net = MLP(S + A, 1, widths, act)
        with torch.no_grad():  # non-zero biases so their gradients / updates are exercised from step 1
            for lin in list(net.layers) + [net.output_layer]:
                lin.b.normal_(0, 0.1)
        opt = torch.optim.Adam(net.parameters(), lr=3e-4)
        tr = MRTrainer(net, opt, num_datapoints=B * K, device="cpu", bt_pool=pool)
        rng = np.random.default_rng(seed)
        batches = [synth_batch(rng, B, T, S, A, ragged=True) for _ in range(K + 1)]
        w0 = [p.detach().numpy().copy() for p in net.parameters()]
        # first-step gradients
        tb = [torch.from_numpy(x) for x in batches[0]]
        losses, accs = [], []
        for k in range(K):
            tb = [torch.from_numpy(x) for x in batches[k]]
            log = tr.train(tb)
            if k == 0:
                g0 = [p.grad.detach().numpy().copy() for p in net.parameters()]
            losses.append(log["training_loss"]); accs.append(log["training_acc"])
        ev = tr.evaluation([torch.from_numpy(x) for x in batches[K]])
        params = list(net.parameters())
        out[f"{name}/meta"] = np.array([S, A, B, T, K, len(widths)] + widths, dtype=np.int64)
        out[f"{name}/act"] = np.array(act)
        out[f"{name}/pool"] = np.array(pool)
        for i, w in enumerate(w0):
            out[f"{name}/w0_{i}"] = w
            out[f"{name}/g0_{i}"] = g0[i]
            out[f"{name}/wK_{i}"] = params[i].detach().numpy().copy()
            out[f"{name}/mK_{i}"] = opt.state[params[i]]["exp_avg"].numpy().copy()
            out[f"{name}/vK_{i}"] = opt.state[params[i]]["exp_avg_sq"].numpy().copy()
        for k, b in enumerate(batches):
            for j in (0, 1, 3, 4, 5, 7, 8):
                out[f"{name}/b{k}_{j}"] = b[j]
        out[f"{name}/loss"] = np.array(losses, dtype=np.float64)
        out[f"{name}/acc"] = np.array(accs, dtype=np.float64)
        out[f"{name}/eval"] = np.array([ev["eval_loss"], ev["eval_acc"]], dtype=np.float64)
        print(name, losses, accs, ev)
    np.savez_compressed(out_path, **out)
    print("wrote", out_path, os.path.getsize(out_path), "bytes")


def load_case(z, name):
    """-> dict(meta fields, w0, g0, wK, mK, vK lists, batches list of 9-tuples, loss, acc, eval)."""
    meta = z[f"{name}/meta"]
    S, A, B, T, K, depth = (int(x) for x in meta[:6])
    widths = [int(x) for x in meta[6:6 + depth]]
    n_par = 2 * (depth + 1)
    ts = np.tile(np.arange(T, dtype=np.int64), (B, 1))
    batches = []
    for k in range(K + 1):
        g = {j: z[f"{name}/b{k}_{j}"] for j in (0, 1, 3, 4, 5, 7, 8)}
        batches.append((g[0], g[1], ts, g[3], g[4], g[5], ts, g[7], g[8]))
    lst = lambda p: [z[f"{name}/{p}_{i}"] for i in range(n_par)]
    return dict(S=S, A=A, B=B, T=T, K=K, widths=widths, act=str(z[f"{name}/act"]), pool=str(z[f"{name}/pool"]),
                w0=lst("w0"), g0=lst("g0"), wK=lst("wK"), mK=lst("mK"), vK=lst("vK"), batches=batches,
                loss=z[f"{name}/loss"], acc=z[f"{name}/acc"], eval=z[f"{name}/eval"])


if __name__ == "__main__":
    here = os.path.dirname(os.path.abspath(__file__))
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(here, "..", "tests", "golden", "mr_train.npz"))
