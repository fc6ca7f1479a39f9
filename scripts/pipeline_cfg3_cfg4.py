"""BASELINE configs[2] and configs[3] end to end inside this package, on synthetic data (no network, random-init weights):

  cfg3  preference transformer trained on synthetic len-100 segment pairs (PTTrainer / fit_pt -> best_model.pt)
        -> load_pt_reward_model -> qlearning_dataset_pt (relabel of the whole dataset) -> modify_reward -> IQL
  cfg4  MR reward MLP trained with snapshots (MRTrainer / fit_mr -> checkpoint_{epoch}.pt) -> qlearning_dataset_mr_ensemble
        (CVaR over the snapshot ensemble) -> IQL on pen-human-shaped data with actor dropout 0.1

Prints one JSON line per pipeline with the wall-clock time of every stage (torch.cuda.synchronize() on both sides).  The
reference runs the same stages as three separate CPU / GPU programs (scripts_pt/run_pt_training.py, scripts_mr/..., iql.py).
Usage: python scripts/pipeline_cfg3_cfg4.py [--rows 200000] [--agents 8] [--iql-steps 500] [--epochs 2]
"""
import argparse
import json
import os
import sys
import tempfile
import time

import numpy as np
import torch
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from iqlpref_b200 import iql as biql  # noqa: E402
from iqlpref_b200.mr_training import MRTrainer, fit_mr  # noqa: E402
from iqlpref_b200.pt_training import PTTrainer, fit_pt  # noqa: E402
from iqlpref_b200.reward_models.pref_transformer import PT  # noqa: E402
from iqlpref_b200.reward_models.q_mlp import MLP as RewardMLP  # noqa: E402


class Env:  # what the relabel functions read from a gym env
    def __init__(self, max_episode_steps):
        self._max_episode_steps = max_episode_steps


def dataset(rng, rows, S, A, ep_len):
    d = {"observations": rng.standard_normal((rows, S)).astype(np.float32),
         "actions": rng.uniform(-1, 1, (rows, A)).astype(np.float32),
         "next_observations": rng.standard_normal((rows, S)).astype(np.float32),
         "rewards": np.zeros(rows, np.float32), "terminals": np.zeros(rows, np.float32)}
    d["terminals"][ep_len - 1::ep_len] = 1
    return d


def pref_batches(rng, n_batches, B, T, S, A, dev):
    out = []
    for _ in range(n_batches):
        b = []
        for _s in range(2):
            b += [torch.from_numpy(rng.standard_normal((B, T, S)).astype(np.float32)).to(dev),
                  torch.from_numpy(rng.uniform(-1, 1, (B, T, A)).astype(np.float32)).to(dev),
                  torch.arange(T, device=dev).repeat(B, 1), torch.ones((B, T), device=dev)]
        y = torch.from_numpy((rng.uniform(size=B) < 0.5).astype(np.float32)).to(dev)
        out.append(b + [torch.stack([y, 1 - y], dim=1)])
    return out


class Clock:
    def __init__(self):
        self.t, self.stages = None, {}

    def __call__(self, name=None):
        torch.cuda.synchronize()
        now = time.perf_counter()
        if name is not None:
            self.stages[name] = round(now - self.t, 4)
        self.t = now


def run_iql(data, S, A, n_agents, steps, dropout, env_name):
    data = dict(data)
    biql.modify_reward(data, env_name, 1, max_episode_steps=1000)  # normalize_reward = 1 (iql.py:363-401; a no-op for pen)
    buf = biql.ReplayBuffer(S, A, data["observations"].shape[0], "cuda")
    buf.load_d4rl_dataset(data)
    eng = biql.BatchedIQL(S, A, n_agents, batch_size=256, precision="bf16", actor_dropout=dropout,
                          hparams=[dict(max_steps=steps)] * n_agents, seeds=list(range(n_agents)))
    for a in range(n_agents):
        torch.manual_seed(a)
        eng.adopt(a, biql.TwinQ(S, A), biql.ValueFunction(S), biql.GaussianPolicy(S, A, 1.0, dropout=dropout), bind=False)
    eng.attach_replay(buf)
    for _ in range(steps):
        eng.train(None)
    torch.cuda.synchronize()
    return eng.losses.cpu().numpy()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=200_000)
    ap.add_argument("--agents", type=int, default=8)
    ap.add_argument("--iql-steps", type=int, default=500)
    ap.add_argument("--epochs", type=int, default=2)
    ap.add_argument("--batches", type=int, default=8)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(0)
    # ---------------------------------------------------------------- cfg3: PT -> relabel -> IQL (antmaze-shaped)
    S, A, T = 29, 8, 100
    with tempfile.TemporaryDirectory() as tmp:
        ck = Clock(); ck()
        torch.manual_seed(0)
        net = PT(S, A, 1000, 256, 256, 4, 0.1, 0.1, 1024, 1, 0.1, 2048, 0.1)
        tr = PTTrainer(net, torch.optim.Adam(net.parameters(), lr=3e-4), device=dev, precision="split_bf16", seed=0)
        train_b, val_b = pref_batches(rng, args.batches, 64, T, S, A, dev), pref_batches(rng, 2, 64, T, S, A, dev)
        ck("setup")
        hist = fit_pt(tr, train_b, val_b, epochs=args.epochs, eval_every=1, checkpoints_path=tmp, criteria_key="loss")
        with open(os.path.join(tmp, "config.yaml"), "w") as f:
            yaml.safe_dump({"num_heads": 4, "intermediate_dim": 1024, "model_eps": 0.1}, f)
        ck("pt_training")
        model = biql.load_pt_reward_model(tmp, "cuda")
        data = dataset(rng, args.rows, S, A, 1000)
        ck("load_model_and_dataset")
        relabelled = biql.qlearning_dataset_pt(Env(1000), model, query_length=T, dataset=data)
        ck("relabel_pt")
        losses = run_iql(relabelled, S, A, args.agents, args.iql_steps, None, "antmaze-large-diverse-v2")
        ck("iql")
        print(json.dumps({"pipeline": "configs[2]: PT training -> relabel -> IQL", "rows": args.rows, "agents": args.agents,
                          "pt_train_steps": args.epochs * args.batches, "iql_steps": args.iql_steps, "seconds": ck.stages,
                          "pt_eval_loss": [round(h["eval_loss"], 5) for h in hist],
                          "reward_range": [float(relabelled["rewards"].min()), float(relabelled["rewards"].max())],
                          "final_losses_agent0": [round(float(x), 5) for x in losses[0]]}), flush=True)
        assert np.isfinite(losses).all()
    # ---------------------------------------------------------------- cfg4: MR snapshots -> CVaR ensemble -> IQL (pen-shaped)
    S, A, T = 45, 24, 100
    with tempfile.TemporaryDirectory() as tmp:
        ck = Clock(); ck()
        torch.manual_seed(1)
        mlp = RewardMLP(S + A, 1, [128, 128, 128], "relu")
        mr = MRTrainer(mlp, torch.optim.Adam(mlp.parameters(), lr=3e-4), 1, device=dev, precision="split_bf16")
        train_b, val_b = pref_batches(rng, args.batches, 256, T, S, A, dev), pref_batches(rng, 2, 256, T, S, A, dev)
        ck("setup")
        epochs = max(args.epochs, 10)
        fit_mr(mr, train_b, val_b, epochs=epochs, eval_every=1, checkpoints_path=tmp, criteria_key="acc", activations="relu")
        ck("mr_training")
        data = dataset(rng, args.rows, S, A, 200)
        ck("dataset")
        relabelled = biql.qlearning_dataset_mr_ensemble(Env(200), tmp, alpha=0.9, burn_in=1, dataset=data)
        ck("relabel_ensemble_cvar")
        losses = run_iql(relabelled, S, A, args.agents, args.iql_steps, 0.1, "pen-human-v1")
        ck("iql")
        print(json.dumps({"pipeline": "configs[3]: MR snapshot ensemble (CVaR 0.9) -> relabel -> IQL (pen-human shape, actor dropout 0.1)",
                          "rows": args.rows, "agents": args.agents, "snapshots": epochs, "mr_train_steps": epochs * args.batches,
                          "iql_steps": args.iql_steps, "seconds": ck.stages,
                          "final_losses_agent0": [round(float(x), 5) for x in losses[0]]}), flush=True)
        assert np.isfinite(losses).all()


if __name__ == "__main__":
    main()
