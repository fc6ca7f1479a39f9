#!/usr/bin/env python
"""Run the UNMODIFIED reference trainer on the B200 (O-ac-cuda of SURVEY.md 8c) and write fixtures + a spread report.

TEST INFRASTRUCTURE; runs on the GPU box, where the reference exists only as the git-ignored copy
`baseline/_ref/` (oracle/make_ref.py):

    gpurun -- 'python scripts/gen_cuda_golden.py'     # -> gpurun_out/cuda_golden/*.npz, spread.json

For cfg1 / cfg2 / cfg4det (the shapes of BASELINE.json configs[0], [1], [3]; same seeds, tables and index
streams as oracle/gen_golden.py) it runs K = 20 steps of `ImplicitQLearning.train` exactly as shipped
(bf16 autocast, iql.py:650, fused Adam :1481-1491) and the fp32 statement of the same step
(algorithms/minari/iql.py:466-488) on `cuda`, stores the per-step losses and per-tensor summaries of the
final parameters (`iql_{cfg}_autocast_cuda.npz`, committed under tests/golden/ afterwards), and measures

  * the reference's own CUDA-autocast <-> fp32 spread (where the bf16 tolerance of the tests comes from),
  * precision="bf16" (the benchmarked tcgen05 path) against both,
  * precision="fp32" against the reference's fp32 run, and the CPU-autocast fixture against the CUDA one.
"""
import json
import os
import sys

import numpy as np
import torch
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.gen_golden import run_reference_steps, summarize  # noqa: E402
from oracle.ref_import import load_reference_iql  # noqa: E402

OUT = os.path.join(ROOT, "gpurun_out", "cuda_golden")
GOLDEN = os.path.join(ROOT, "tests", "golden")
CFGS = {"cfg1": (17, 6, False), "cfg2": (29, 8, False), "cfg4det": (45, 24, True)}
HP = dict(discount=0.99, tau=0.005, beta=3.0, iql_tau=0.7, vf_lr=3e-4, qf_lr=3e-4, actor_lr=3e-4, max_steps=1000)
K, N_ROWS, SEED = 20, 1024, 7


def dtype_probe(ref):
    """which intermediate tensors of train() are bf16 under CUDA autocast."""
    S, A = 17, 6
    torch.manual_seed(0)
    qf, vf, ac = ref.TwinQ(S, A).cuda(), ref.ValueFunction(S).cuda(), ref.GaussianPolicy(S, A, 1.0).cuda()
    obs, act = torch.randn(8, S, device="cuda"), torch.rand(8, A, device="cuda")
    out = {}
    with torch.amp.autocast("cuda", dtype=torch.bfloat16):
        v = vf(obs)
        tq = qf(obs, act)
        adv = tq - v
        out["v"], out["target_q"], out["adv"] = str(v.dtype), str(tq.dtype), str(adv.dtype)
        out["adv**2"] = str((adv ** 2).dtype)
        out["v_loss"] = str(ref.asymmetric_l2_loss(adv, 0.7).dtype)
        out["beta*adv"] = str((3.0 * adv.detach()).dtype)
        out["exp_adv"] = str(torch.exp(3.0 * adv.detach()).dtype)
        targets = torch.randn(8, device="cuda") + (1.0 - torch.zeros(8, device="cuda")) * 0.99 * v.detach()
        out["targets"] = str(targets.dtype)
        out["mse"] = str(torch.nn.functional.mse_loss(tq, targets).dtype)
        dist = ac(obs)
        out["policy_mean"] = str(dist.mean.dtype)
        out["log_prob"] = str(dist.log_prob(act).dtype)
        loss = ref.asymmetric_l2_loss(adv, 0.7)
    loss.backward()
    out["grad_W"] = str(vf.v.net[0].weight.grad.dtype)
    return out


def tensor_err(a, b):
    """max |a-b| normalised by max |b| over the strided samples (the fixtures' parameter criterion)."""
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-12))


def loss_err(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return [float(x) for x in (np.abs(a - b) / np.maximum(np.abs(b), 1e-6)).max(axis=0)]


def our_run(cfg, precision, hp):
    from iqlpref_b200 import iql as biql
    S, A, det = CFGS[cfg]
    torch.manual_seed(SEED)
    qf, vf = biql.TwinQ(S, A), biql.ValueFunction(S)
    actor = (biql.DeterministicPolicy if det else biql.GaussianPolicy)(S, A, 1.0)
    eng = biql.BatchedIQL(S, A, 1, batch_size=256, deterministic=det, hparams=[hp], precision=precision)
    eng.adopt(0, qf, vf, actor)
    rng = np.random.default_rng(SEED + 1000)
    tab = {"obs": rng.standard_normal((N_ROWS, S)).astype(np.float32),
           "act": rng.uniform(-1, 1, (N_ROWS, A)).astype(np.float32),
           "rew": (rng.standard_normal(N_ROWS) * 0.5 + 0.5).astype(np.float32),
           "next_obs": rng.standard_normal((N_ROWS, S)).astype(np.float32),
           "done": (rng.uniform(size=N_ROWS) < 0.05).astype(np.float32)}
    idx = rng.integers(0, N_ROWS, size=(K, 256)).astype(np.int64)
    losses = np.zeros((K, 3))
    for k in range(K):
        i = torch.from_numpy(idx[k])
        rows = biql.pack_rows(torch.from_numpy(tab["obs"])[i], torch.from_numpy(tab["act"])[i],
                              torch.from_numpy(tab["rew"])[i], torch.from_numpy(tab["next_obs"])[i],
                              torch.from_numpy(tab["done"])[i], eng.R)
        eng.train(rows.cuda().unsqueeze(0))
        losses[k] = eng.losses[0].cpu().numpy()
    final = {}
    for nm, mod in (("qf", qf), ("vf", vf), ("actor", actor)):
        for key, val in mod.state_dict().items():
            final[f"{nm}.{key}"] = val.detach().cpu().numpy().copy()
    e = eng.export(0)
    for net in ("q1", "q2"):
        for i in range(3):
            final[f"q_target.{net}.net.{2 * i}.weight"] = e[f"qt{net[1]}.w{i}"]
            final[f"q_target.{net}.net.{2 * i}.bias"] = e[f"qt{net[1]}.b{i}"]
    return losses, final


def compare(name, la, fa, lb, fb, report):
    per_tensor = {k: tensor_err(fa[k].reshape(-1)[::97], fb[k].reshape(-1)[::97]) for k in fb if k in fa}
    report[name] = {"loss_max_rel": loss_err(la, lb), "param_max_norm_err": max(per_tensor.values()),
                    "worst_tensor": max(per_tensor, key=per_tensor.get)}


def main():
    os.makedirs(OUT, exist_ok=True)
    ref = load_reference_iql()
    torch.backends.cudnn.benchmark = True
    report = {"torch": torch.__version__, "device": torch.cuda.get_device_name(0), "dtype_probe": dtype_probe(ref),
              "tf32_matmul_allowed": bool(torch.backends.cuda.matmul.allow_tf32), "K": K, "cfgs": {}}
    hp2 = dict(HP, beta=10.0, iql_tau=0.9)
    for cfg, (S, A, det) in CFGS.items():
        hp = hp2 if cfg == "cfg2" else HP
        rep = report["cfgs"][cfg] = {}
        runs = {}
        for mode in ("autocast", "fp32"):
            _, idx, init, final, losses, lr = run_reference_steps(ref, S, A, 256, 256, K, N_ROWS, SEED, mode, det, hp,
                                                                  device="cuda")
            runs[mode] = (losses, final)
            if mode == "autocast":
                # twice: is the reference's CUDA run reproducible bit for bit?
                _, _, _, final2, losses2, _ = run_reference_steps(ref, S, A, 256, 256, K, N_ROWS, SEED, mode, det, hp,
                                                                  device="cuda")
                rep["ref_autocast_cuda_reproducible"] = bool(np.array_equal(losses, losses2))
            np.savez(os.path.join(OUT, f"iql_{cfg}_{mode}_cuda.npz"), idx=idx, losses=losses, lr_final=np.array(lr),
                     hp=np.array(yaml.safe_dump(hp)), seed=np.array(SEED), n_rows=np.array(N_ROWS),
                     **{f"init.{k}": v for k, v in summarize(init).items()},
                     **{f"final.{k}": v for k, v in summarize(final).items()})
        compare("ref_autocast_cuda_vs_ref_fp32_cuda", *runs["autocast"], *runs["fp32"], rep)
        for prec in ("bf16", "fp32"):
            try:
                lo, fo = our_run(cfg, prec, hp)
            except Exception as e:  # noqa: BLE001
                rep[f"ours_{prec}_error"] = repr(e)
                continue
            compare(f"ours_{prec}_vs_ref_autocast_cuda", lo, fo, *runs["autocast"], rep)
            compare(f"ours_{prec}_vs_ref_fp32_cuda", lo, fo, *runs["fp32"], rep)
            rep[f"ours_{prec}_losses_first3"] = lo[:3].tolist()
        rep["ref_autocast_cuda_losses_first3"] = runs["autocast"][0][:3].tolist()
        rep["ref_fp32_cuda_losses_first3"] = runs["fp32"][0][:3].tolist()
        for tag in ("autocast", "fp32"):  # committed CPU fixtures against the CUDA runs
            d = np.load(os.path.join(GOLDEN, f"iql_{cfg}_{tag}.npz"))
            per = {k: tensor_err(d[f"final.{k}#sample"], runs[tag][1][k].reshape(-1)[::97]) for k in runs[tag][1]}
            rep[f"cpu_{tag}_fixture_vs_cuda_{tag}"] = {"loss_max_rel": loss_err(d["losses"], runs[tag][0]),
                                                       "param_max_norm_err": max(per.values())}
    with open(os.path.join(OUT, "spread.json"), "w") as f:
        json.dump(report, f, indent=1)
    print(json.dumps(report, indent=1))


if __name__ == "__main__":
    main()
