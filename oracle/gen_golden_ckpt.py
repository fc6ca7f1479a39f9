"""Generate the checkpoint fixtures by running the UNMODIFIED reference in this container.

TEST INFRASTRUCTURE.  Run here (needs /root/reference):  python -m oracle.gen_golden_ckpt

  tests/golden/iql_small_ckpt.pt          what the reference's training loop writes as checkpoint_{t}.pt
                                          (iql.py:1557-1561 -> ImplicitQLearning.state_dict :664-674) after 4 updates
  tests/golden/iql_small_ckpt_dropout.pt  the same file for an actor built with actor_dropout=0.1 (Linear layers at
                                          net.net.0 / 3 / 6), after 2 updates
  tests/golden/iql_small_ckpt_resume.npz  the replay rows / indices of those updates, and what the reference does
                                          when a FRESH trainer loads that file (iql.py:1519-1521, :676-688) and
                                          runs 3 more updates: losses, parameters, target, Adam moments, actor lr
"""
import os
import sys
import warnings

import numpy as np
import torch
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.gen_golden import OUT, flat_modules  # noqa: E402
from oracle.ref_import import load_reference_iql  # noqa: E402

warnings.filterwarnings("ignore")
S, A, H, B, K1, K2, N_ROWS = 17, 6, 64, 64, 4, 3, 300
HP = dict(discount=0.99, tau=0.005, beta=3.0, iql_tau=0.7, vf_lr=3e-4, qf_lr=3e-4, actor_lr=3e-4, max_steps=50)


def make_trainer(ref, seed, dropout=None):
    """Networks and optimizers as iql.py:1467-1491 builds them (non-fused Adam: CPU)."""
    torch.manual_seed(seed)
    qf, vf, actor = ref.TwinQ(S, A, H), ref.ValueFunction(S, H), ref.GaussianPolicy(S, A, 1.0, H, dropout=dropout)
    v_opt = torch.optim.Adam(vf.parameters(), lr=HP["vf_lr"])
    q_opt = torch.optim.Adam(qf.parameters(), lr=HP["qf_lr"])
    a_opt = torch.optim.Adam(actor.parameters(), lr=HP["actor_lr"])
    return ref.ImplicitQLearning(1.0, actor, a_opt, qf, q_opt, vf, v_opt, iql_tau=HP["iql_tau"], beta=HP["beta"],
                                 max_steps=HP["max_steps"], discount=HP["discount"], tau=HP["tau"], device="cpu")


def step_fp32(tr, table, i):
    """One update in the fp32 statement of train() (algorithms/minari/iql.py:466-488)."""
    i = torch.from_numpy(i)
    obs, act, nobs = (torch.from_numpy(table[k])[i] for k in ("obs", "act", "next_obs"))
    rew, done = torch.from_numpy(table["rew"])[i], torch.from_numpy(table["done"])[i]
    tr.total_it += 1
    ld = {}
    with torch.no_grad():
        next_v = tr.vf(nobs)
    adv = tr._update_v(obs, act, ld)
    tr._update_q(next_v, obs, act, rew, done, ld)
    tr._update_policy(adv, obs, act, ld)
    return ld["value_loss"], ld["q_loss"], ld["actor_loss"]


def main():
    ref = load_reference_iql()
    rng = np.random.default_rng(77)
    table = {"obs": rng.standard_normal((N_ROWS, S)).astype(np.float32),
             "act": rng.uniform(-1, 1, (N_ROWS, A)).astype(np.float32),
             "rew": (rng.standard_normal(N_ROWS) * 0.5 + 0.5).astype(np.float32),
             "next_obs": rng.standard_normal((N_ROWS, S)).astype(np.float32),
             "done": (rng.uniform(size=N_ROWS) < 0.05).astype(np.float32)}
    idx = rng.integers(0, N_ROWS, size=(K1 + K2, B)).astype(np.int64)
    tr = make_trainer(ref, seed=3)
    init = flat_modules(tr.qf, tr.vf, tr.actor)
    losses = np.zeros((K1 + K2, 3))
    for k in range(K1):
        losses[k] = step_fp32(tr, table, idx[k])
    path = os.path.join(OUT, "iql_small_ckpt.pt")
    torch.save(tr.state_dict(), path)  # iql.py:1557-1561
    # a fresh process resumes from the file (iql.py:1519-1521)
    tr2 = make_trainer(ref, seed=99)
    tr2.load_state_dict(torch.load(path, weights_only=False))
    for k in range(K1, K1 + K2):
        losses[k] = step_fp32(tr2, table, idx[k])
    final = flat_modules(tr2.qf, tr2.vf, tr2.actor)
    for k_, v_ in tr2.q_target.state_dict().items():
        final[f"q_target.{k_}"] = v_.detach().numpy().copy()
    sd2 = tr2.state_dict()
    for opt in ("q_optimizer", "v_optimizer", "actor_optimizer"):
        for i, st in sd2[opt]["state"].items():
            final[f"{opt}.{i}.exp_avg"] = st["exp_avg"].numpy().copy()
            final[f"{opt}.{i}.exp_avg_sq"] = st["exp_avg_sq"].numpy().copy()
            final[f"{opt}.{i}.step"] = np.array(float(st["step"]))
    np.savez(os.path.join(OUT, "iql_small_ckpt_resume.npz"), **{f"table.{k}": v for k, v in table.items()}, idx=idx,
             losses=losses, k_saved=np.array(K1), hp=np.array(yaml.safe_dump(HP)),
             lr_saved=np.array(torch.load(path, weights_only=False)["actor_optimizer"]["param_groups"][0]["lr"]),
             lr_final=np.array(sd2["actor_optimizer"]["param_groups"][0]["lr"]), total_it=np.array(sd2["total_it"]),
             sched_last_epoch=np.array(sd2["actor_lr_schedule"]["last_epoch"]),
             **{f"init.{k}": v for k, v in init.items()}, **{f"final.{k}": v for k, v in final.items()})
    # an actor with nn.Dropout (pen / door configs, actor_dropout: 0.1): its Linear layers sit at net.net.0 / 3 / 6
    # (iql.py:436-437).  Two updates (the masks come from torch's CPU generator: only the FORMAT is pinned here).
    tr3 = make_trainer(ref, seed=5, dropout=0.1)
    for k in range(2):
        step_fp32(tr3, table, idx[k])
    torch.save(tr3.state_dict(), os.path.join(OUT, "iql_small_ckpt_dropout.pt"))
    for f in ("iql_small_ckpt.pt", "iql_small_ckpt_resume.npz", "iql_small_ckpt_dropout.pt"):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
