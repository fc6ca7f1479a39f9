"""Generate tests/golden/*.npz by running the UNMODIFIED reference in this container.

TEST INFRASTRUCTURE.  Run here (needs /root/reference):  python -m oracle.gen_golden
The fixtures travel to the GPU box; /root/reference does not.
"""
import os
import sys
import tempfile
import warnings

import numpy as np
import torch
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.ref_import import load_reference_iql  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
warnings.filterwarnings("ignore")


class FakeEnv:
    def __init__(self, max_episode_steps):
        self._max_episode_steps = max_episode_steps


def synth_dataset(rng, n, s, a, ep_len, p_term=0.0, with_timeouts=True):
    d = {
        "observations": rng.standard_normal((n, s)).astype(np.float32),
        "actions": rng.uniform(-1, 1, (n, a)).astype(np.float32),
        "rewards": (rng.standard_normal(n) * 0.5 + 0.5).astype(np.float32),
        "terminals": (rng.uniform(size=n) < p_term),
    }
    if with_timeouts:
        t = np.zeros(n, dtype=bool)
        t[ep_len - 1::ep_len] = True
        d["timeouts"] = t
    return d


def flat_modules(qf, vf, actor):
    d = {}
    for name, mod in (("qf", qf), ("vf", vf), ("actor", actor)):
        for k, v in mod.state_dict().items():
            d[f"{name}.{k}"] = v.detach().cpu().numpy().copy()
    return d


def run_reference_steps(ref, S, A, H, B, K, n_rows, seed, mode, deterministic, hp, device="cpu"):
    """Build reference nets exactly like iql.py:1467-1491 and run K steps on fixed indices.  `device="cuda"`
    (scripts/gen_cuda_golden.py, on the GPU box) follows the reference's CUDA set-up: modules moved to the device
    after their CPU initialisation, `Adam(fused=True)` (iql.py:1470-1491)."""
    rng = np.random.default_rng(seed + 1000)
    table = {
        "obs": rng.standard_normal((n_rows, S)).astype(np.float32),
        "act": rng.uniform(-1, 1, (n_rows, A)).astype(np.float32),
        "rew": (rng.standard_normal(n_rows) * 0.5 + 0.5).astype(np.float32),
        "next_obs": rng.standard_normal((n_rows, S)).astype(np.float32),
        "done": (rng.uniform(size=n_rows) < 0.05).astype(np.float32),
    }
    idx = rng.integers(0, n_rows, size=(K, B)).astype(np.int64)
    torch.manual_seed(seed)
    qf = ref.TwinQ(S, A, H)
    vf = ref.ValueFunction(S, H)
    actor = (ref.DeterministicPolicy if deterministic else ref.GaussianPolicy)(S, A, 1.0, H)
    init = flat_modules(qf, vf, actor)
    qf, vf, actor = qf.to(device), vf.to(device), actor.to(device)
    kw = {"fused": True} if str(device).startswith("cuda") else {}
    v_opt = torch.optim.Adam(vf.parameters(), lr=hp["vf_lr"], **kw)
    q_opt = torch.optim.Adam(qf.parameters(), lr=hp["qf_lr"], **kw)
    a_opt = torch.optim.Adam(actor.parameters(), lr=hp["actor_lr"], **kw)
    tr = ref.ImplicitQLearning(1.0, actor, a_opt, qf, q_opt, vf, v_opt, iql_tau=hp["iql_tau"], beta=hp["beta"],
                               max_steps=hp["max_steps"], discount=hp["discount"], tau=hp["tau"], device=str(device))
    losses = np.zeros((K, 3), dtype=np.float64)
    dtab = {k: torch.from_numpy(v).to(device) for k, v in table.items()}
    for k in range(K):
        i = torch.from_numpy(idx[k]).to(device)
        batch = [dtab["obs"][i], dtab["act"][i], dtab["rew"][i, None], dtab["next_obs"][i], dtab["done"][i, None]]
        if mode == "autocast":
            ld = tr.train(batch)  # as shipped: bf16 autocast, iql.py:650
        else:  # fp32 statement of the same step (algorithms/minari/iql.py:466-488)
            obs, act, rew, nobs, done = batch
            tr.total_it += 1
            ld = {}
            with torch.no_grad():
                next_v = tr.vf(nobs)
            adv = tr._update_v(obs, act, ld)
            tr._update_q(next_v, obs, act, rew.squeeze(-1), done.squeeze(-1), ld)
            tr._update_policy(adv, obs, act, ld)
        losses[k] = (ld["value_loss"], ld["q_loss"], ld["actor_loss"])
    final = flat_modules(qf, vf, actor)
    for k_, v_ in tr.q_target.state_dict().items():
        final[f"q_target.{k_}"] = v_.detach().cpu().numpy().copy()
    lr_final = a_opt.param_groups[0]["lr"]
    return table, idx, init, final, losses, lr_final


def summarize(d):
    """Per-tensor (sum, l2, strided sample) for fixtures too large to store whole."""
    out = {}
    for k, v in d.items():
        f = v.reshape(-1).astype(np.float64)
        out[k + "#sum"] = np.array(f.sum())
        out[k + "#l2"] = np.array(np.sqrt((f * f).sum()))
        out[k + "#sample"] = v.reshape(-1)[::97].copy()
    return out


def gen_iql(ref):
    hp = dict(discount=0.99, tau=0.005, beta=3.0, iql_tau=0.7, vf_lr=3e-4, qf_lr=3e-4, actor_lr=3e-4, max_steps=1000)
    # small, full tensors stored
    for name, det in (("gauss", False), ("det", True)):
        table, idx, init, final, losses, lr = run_reference_steps(ref, 17, 6, 64, 64, 12, 300, 3, "fp32", det, hp)
        np.savez(os.path.join(OUT, f"iql_small_{name}_fp32.npz"), **{f"table.{k}": v for k, v in table.items()},
                 idx=idx, losses=losses, lr_final=np.array(lr), hp=np.array(yaml.safe_dump(hp)),
                 **{f"init.{k}": v for k, v in init.items()}, **{f"final.{k}": v for k, v in final.items()})
    # full-size nets (2x256), summaries only; init is regenerated from the seed by the host mirror
    hp2 = dict(hp, beta=10.0, iql_tau=0.9)
    for cfg, (S, A, det, h) in {"cfg1": (17, 6, False, hp), "cfg2": (29, 8, False, hp2), "cfg4det": (45, 24, True, hp)}.items():
        for mode in ("fp32", "autocast"):
            table, idx, init, final, losses, lr = run_reference_steps(ref, S, A, 256, 256, 20, 1024, 7, mode, det, h)
            np.savez(os.path.join(OUT, f"iql_{cfg}_{mode}.npz"), idx=idx, losses=losses, lr_final=np.array(lr),
                     hp=np.array(yaml.safe_dump(h)), seed=np.array(7), n_rows=np.array(1024),
                     **{f"init.{k}": v for k, v in summarize(init).items()},
                     **{f"final.{k}": v for k, v in summarize(final).items()})


def save_mr_ckpt(ref, path, net):
    torch.save({"net": net.state_dict(), "optimizer": {}}, path)


def gen_relabel(ref):
    rng = np.random.default_rng(11)
    S, A = 29, 8
    out = {}
    # --- RewardMLP forward for several activations
    x = rng.standard_normal((200, S + A)).astype(np.float32)
    out["mlp.x"] = x
    for act in ("relu", "tanh", "swish", "softplus", "leaky_relu", "rbf", "sin", "cos", "linear"):
        torch.manual_seed(5)
        net = ref.RewardMLP(S + A, 1, [32, 48, 32], act)
        with torch.no_grad():
            net.layers[0].b.normal_(); net.output_layer.b.normal_()
        for i, prm in enumerate(net.parameters()):
            out[f"mlp.{act}.w{i}"] = prm.detach().numpy().copy()
        with torch.inference_mode():
            out[f"mlp.{act}.y"] = net(torch.from_numpy(x)).numpy().copy()
    # --- datasets
    ds_t = synth_dataset(rng, 400, S, A, ep_len=50, p_term=0.02, with_timeouts=True)
    ds_n = synth_dataset(rng, 400, S, A, ep_len=50, p_term=0.02, with_timeouts=False)
    for tag, ds in (("ds_t", ds_t), ("ds_n", ds_n)):
        for k, v in ds.items():
            out[f"{tag}.{k}"] = v
    env = FakeEnv(60)
    # --- qlearning_dataset_mr
    torch.manual_seed(6)
    mr = ref.RewardMLP(S + A, 1, [32, 32, 32], "relu")
    for i, prm in enumerate(mr.parameters()):
        out[f"mr.w{i}"] = prm.detach().numpy().copy()
    for tag, ds in (("ds_t", ds_t), ("ds_n", ds_n)):
        for toe in (False, True):
            r = ref.qlearning_dataset_mr(env, mr, dataset=dict(ds), terminate_on_end=toe)
            for k, v in r.items():
                out[f"mr.{tag}.toe{int(toe)}.{k}"] = v
    # --- MR snapshot ensemble
    with tempfile.TemporaryDirectory() as td:
        with open(os.path.join(td, "config.yaml"), "w") as f:
            yaml.safe_dump({"activations": "tanh"}, f)
        for e in range(0, 60, 5):
            torch.manual_seed(100 + e)
            net = ref.RewardMLP(S + A, 1, [32, 32], "tanh")
            sd = net.state_dict()
            if e % 10 == 0:  # some snapshots saved by a torch.compile'd model
                sd = {"_orig_mod." + k: v for k, v in sd.items()}
            torch.save({"net": sd}, os.path.join(td, f"checkpoint_{e}.pt"))
            for i, prm in enumerate(net.parameters()):
                out[f"ens.e{e}.w{i}"] = prm.detach().numpy().copy()
        torch.save({"net": net.state_dict()}, os.path.join(td, "best_model.pt"))
        for alpha in (0.0, 0.5, 0.9, 0.95):
            for burn in (0, 10):
                r = ref.qlearning_dataset_mr_ensemble(env, td, alpha=alpha, burn_in=burn, dataset=dict(ds_t))
                out[f"ens.a{alpha}.b{burn}.rewards"] = r["rewards"]
    # --- BNN posterior (tuple-of-arrays format, bayes_net.py:92-98)
    with tempfile.TemporaryDirectory() as td:
        for c in range(2):
            d = os.path.join(td, "sampling_f", f"chain_{c}", "sampled_weights")
            os.makedirs(d)
            samples = []
            for k in range(20):
                ws = []
                dims = [S + A, 16, 16, 16, 1]
                for l in range(4):
                    ws.append(rng.standard_normal((dims[l], dims[l + 1])).astype(np.float32))
                    ws.append((rng.standard_normal(dims[l + 1]) * 0.1).astype(np.float32))
                samples.append(tuple(ws))
                for i, w in enumerate(ws):
                    out[f"bnn.c{c}.k{k}.w{i}"] = w
            torch.save({"sampled_weights": samples}, os.path.join(d, "sampled_weights_0000000"))
        for alpha, ns in ((0.9, 500), (0.9, 16), (0.0, 40), (0.75, 40)):
            r = ref.qlearning_dataset_bnn(env, td, alpha=alpha, n_samples=ns, device="cpu", dataset=dict(ds_t))
            out[f"bnn.a{alpha}.n{ns}.rewards"] = r["rewards"]
    # --- PT
    for L in (1, 2):
        torch.manual_seed(20 + L)
        pt = ref.RewardPT(state_dim=S, action_dim=A, max_episode_steps=60, embd_dim=32, pref_attn_embd_dim=32,
                          num_heads=4, intermediate_dim=64, num_layers=L, max_pos=64, eps=0.1)
        pt.eval()
        with torch.no_grad():  # make LN affine / biases non-trivial
            for n_, p_ in pt.named_parameters():
                if "layer_norm" in n_ or n_.endswith("bias"):
                    p_.add_(torch.randn_like(p_) * 0.2)
        for k, v in pt.state_dict().items():
            if "causal_bias" not in k:
                out[f"pt{L}.{k}"] = v.numpy().copy()
        sts = rng.standard_normal((6, 10, S)).astype(np.float32)
        acts = rng.uniform(-1, 1, (6, 10, A)).astype(np.float32)
        ts = np.tile(np.arange(10), (6, 1)).astype(np.int64)
        am = np.ones((6, 10), np.float32)
        for j, padn in enumerate((0, 3, 9, 0, 5, 1)):
            sts[j, :padn] = 0; acts[j, :padn] = 0; am[j, :padn] = 0
            ts[j, :padn] = 0; ts[j, padn:] = np.arange(10 - padn)
        with torch.inference_mode():
            res, _ = pt(torch.from_numpy(sts), torch.from_numpy(acts), torch.from_numpy(ts), torch.from_numpy(am))
        out[f"pt{L}.fw.sts"], out[f"pt{L}.fw.acts"], out[f"pt{L}.fw.ts"], out[f"pt{L}.fw.am"] = sts, acts, ts, am
        out[f"pt{L}.fw.value"] = res["value"][:, 0, :, 0].numpy().copy()
        out[f"pt{L}.fw.weighted_sum"] = res["weighted_sum"][:, :, 0].numpy().copy()
        for tag, ds in (("ds_t", ds_t), ("ds_n", ds_n)):
            r = ref.qlearning_dataset_pt(env, pt, query_length=10, dataset=dict(ds))
            out[f"pt{L}.{tag}.rewards"] = r["rewards"]
    # --- modify_reward
    for envname in ("antmaze-large-diverse-v2", "halfcheetah-medium-v2"):
        for nr in range(1, 9):
            d = {"rewards": ds_t["rewards"].copy(), "terminals": ds_t["terminals"].copy()}
            ref.modify_reward(d, envname, nr, max_episode_steps=50)
            out[f"modrew.{envname}.{nr}"] = d["rewards"]
    # --- empirical_cvar
    smp = rng.standard_normal(37).astype(np.float32)
    out["cvar.samples"] = smp
    out["cvar.values"] = np.array([ref.empirical_cvar(smp, a) for a in (0.0, 0.5, 0.9, 0.95, 0.99)])
    np.savez(os.path.join(OUT, "relabel.npz"), **out)


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    ref = load_reference_iql()
    gen_iql(ref)
    gen_relabel(ref)
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))
