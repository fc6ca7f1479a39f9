"""Fixture for the CVaR tail membership against the REFERENCE's fp32 predictions.

TEST INFRASTRUCTURE.  Run here (needs /root/reference):  python -m oracle.gen_golden_cvar

For S in {40, 981} members of `RewardMLP(69, 1, [64]*3, "relu")` (the pen-human BNN / snapshot shapes of SURVEY 8d
cfg4; weights drawn from `members_from_seed`, which the test repeats) the unmodified reference module
(optbnn/bnn/nets/mlp.py:73-111, fp32 on the CPU exactly as qlearning_dataset_bnn / _mr_ensemble evaluate it,
iql.py:982-1011, 1162-1187) predicts n = 192 rows; `tests/golden/cvar_ref.npz` keeps those `all_preds` matrices.  The
test recomputes the reference's tail sets from them (`np.argpartition(all_preds, kth, axis=0)[:n_tail]`, iql.py:1009-1011)
and compares them with the tail membership the CUDA path selects from ITS OWN predictions in the default precision.
"""
import os
import sys
import warnings

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.ref_import import load_reference_iql  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
D_IN, WIDTH, DEPTH, N_ROWS = 69, 64, 3, 192
warnings.filterwarnings("ignore")


def members_from_seed(S, seed=0):
    """[W0, b0, ..., Wout, bout] per member, W stored (n_in, n_out) like the BNN tuples (bayes_net.py:92-98).  A shared
    mean network plus a per-member perturbation, so that member predictions of one row are close together - the regime
    in which tail membership is sensitive to the arithmetic."""
    rng = np.random.default_rng(seed)
    dims = [D_IN] + [WIDTH] * DEPTH + [1]
    base = [(rng.standard_normal((dims[l], dims[l + 1])).astype(np.float32),
             (rng.standard_normal(dims[l + 1]) * 0.1).astype(np.float32)) for l in range(DEPTH + 1)]
    out = []
    for _ in range(S):
        m = []
        for w, b in base:
            m += [(w + 0.15 * rng.standard_normal(w.shape)).astype(np.float32),
                  (b + 0.05 * rng.standard_normal(b.shape)).astype(np.float32)]
        out.append(m)
    return out


def rows_from_seed(seed=1):
    return np.random.default_rng(seed).standard_normal((N_ROWS, D_IN)).astype(np.float32)


def main():
    ref = load_reference_iql()
    x = torch.from_numpy(rows_from_seed())
    out = {}
    for S in (40, 981):
        net = ref.RewardMLP(D_IN, 1, [WIDTH] * DEPTH, "relu")
        preds = np.zeros((S, N_ROWS), np.float32)
        for s, m in enumerate(members_from_seed(S, seed=S)):
            with torch.no_grad():
                for prm, w in zip(net.parameters(), m):  # param.copy_ per sample, iql.py:986-991
                    prm.copy_(torch.from_numpy(w))
            with torch.inference_mode():
                preds[s] = net(x).squeeze(-1).numpy()
        out[f"preds.S{S}"] = preds
    np.savez_compressed(os.path.join(OUT, "cvar_ref.npz"), **out)
    print("cvar_ref.npz", os.path.getsize(os.path.join(OUT, "cvar_ref.npz")))


if __name__ == "__main__":
    main()
