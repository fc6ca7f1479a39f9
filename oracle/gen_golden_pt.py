"""Fixtures of the preference-transformer training step, written by the UNMODIFIED reference `PTTrainer`
(gp_reward-priors/optbnn/training/training.py:308-416) on its own `PT` (optbnn/bnn/nets/pref_trans.py) with every
dropout at 0 (its masks come from torch's generator and cannot be mirrored; dropout parity is tested by mask injection).

TEST INFRASTRUCTURE.  Run here (needs /root/reference):  python -m oracle.gen_golden_pt  ->  tests/golden/pt_train.npz

Per case: initial state_dict, K + 1 batches, first-step gradients of every parameter, per-step loss / accuracy,
`evaluation()` on the extra batch, final weights and Adam moments after K steps.
"""
import os
import sys
import warnings

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.pt_train_oracle import synth_pt_batch  # noqa: E402
from oracle.ref_import import load_reference_iql  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
warnings.filterwarnings("ignore")

# name -> (S, A, T, B, E, P, heads, I, L, bt_pool, K)
CASES = {
    "l1_mean": (7, 3, 6, 5, 32, 32, 4, 64, 1, "mean", 3),
    "l2_sum": (5, 2, 8, 4, 32, 16, 2, 48, 2, "sum", 3),
}
EPS = 0.1
BATCH_KEYS = ("s1", "a1", "t1", "m1", "s2", "a2", "t2", "m2", "labels")


def build(ref, case, seed):
    S, A, T, B, E, P, heads, I, L, pool, K = CASES[case]
    torch.manual_seed(seed)
    net = ref.RewardPT(state_dim=S, action_dim=A, max_episode_steps=T + 4, embd_dim=E, pref_attn_embd_dim=P, num_heads=heads,
                       attn_dropout=0.0, resid_dropout=0.0, intermediate_dim=I, num_layers=L, embd_dropout=0.0,
                       max_pos=2 * T + 4, eps=EPS)
    for m in net.modules():  # GPT2Model does not forward resid_dropout to its blocks (pref_trans.py:203-213): force 0
        if isinstance(m, torch.nn.Dropout):
            m.p = 0.0
    with torch.no_grad():
        for n_, p_ in net.named_parameters():
            if "layer_norm" in n_ or n_.endswith("bias"):
                p_.add_(torch.randn_like(p_) * 0.2)
    return net


def state(net):
    return {k: v.detach().numpy().copy() for k, v in net.state_dict().items() if "causal_bias" not in k}


def main():
    ref = load_reference_iql()
    from optbnn.training.training import PTTrainer
    out = {}
    for ci, case in enumerate(sorted(CASES)):
        S, A, T, B, E, P, heads, I, L, pool, K = CASES[case]
        rng = np.random.default_rng(100 + ci)
        net = build(ref, case, 7 + ci)
        for k, v in state(net).items():
            out[f"{case}.w0.{k}"] = v
        opt = torch.optim.Adam(net.parameters(), lr=3e-4)
        tr = PTTrainer(net, opt, device="cpu", bt_pool=pool)
        losses, accs = [], []
        for k in range(K + 1):
            batch = synth_pt_batch(rng, B, T, S, A)
            for name, arr in zip(BATCH_KEYS, batch):
                out[f"{case}.batch{k}.{name}"] = arr
            tb = [torch.from_numpy(x) for x in batch]
            if k == K:
                ev = tr.evaluation(tb)
                out[f"{case}.eval"] = np.array([ev["eval_loss"], ev["eval_acc"]])
                break
            if k == 0:  # gradients of the first step (train() zeroes them at its start, they survive its end)
                log = tr.train(tb)
                for n_, p_ in net.named_parameters():
                    out[f"{case}.g0.{n_}"] = (p_.grad.detach().numpy().copy() if p_.grad is not None
                                               else np.zeros(tuple(p_.shape), np.float32))
            else:
                log = tr.train(tb)
            losses.append(log["training_loss"]); accs.append(log["training_acc"])
        out[f"{case}.loss"], out[f"{case}.acc"] = np.array(losses), np.array(accs)
        for k, v in state(net).items():
            out[f"{case}.wK.{k}"] = v
        sd = opt.state_dict()["state"]
        for i, (n_, _) in enumerate(net.named_parameters()):
            if i in sd:
                out[f"{case}.mK.{n_}"] = sd[i]["exp_avg"].numpy().copy()
                out[f"{case}.vK.{n_}"] = sd[i]["exp_avg_sq"].numpy().copy()
    np.savez_compressed(os.path.join(OUT, "pt_train.npz"), **out)
    print("pt_train.npz", os.path.getsize(os.path.join(OUT, "pt_train.npz")))


def load_case(d, case):
    S, A, T, B, E, P, heads, I, L, pool, K = CASES[case]
    pre = case + "."
    get = lambda tag: {k[len(pre + tag) + 1:]: d[k] for k in d.files if k.startswith(pre + tag + ".")}
    batches = [[d[f"{case}.batch{k}.{n}"] for n in BATCH_KEYS] for k in range(K + 1)]
    return dict(S=S, A=A, T=T, B=B, E=E, P=P, heads=heads, I=I, L=L, pool=pool, K=K, eps=EPS, w0=get("w0"), g0=get("g0"),
                wK=get("wK"), mK=get("mK"), vK=get("vK"), batches=batches, loss=d[f"{case}.loss"], acc=d[f"{case}.acc"],
                eval=d[f"{case}.eval"])


if __name__ == "__main__":
    main()
