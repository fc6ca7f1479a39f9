"""The CPU oracle against the imported, UNMODIFIED reference on FRESH inputs (not the committed fixtures): same seeds,
same batches, K steps of `ImplicitQLearning.train` as shipped (bf16 autocast, iql.py:650) and of its fp32 statement
(algorithms/minari/iql.py:466-488).  Runs wherever the reference can be imported (`/root/reference` in the authoring
container, `baseline/_ref` on the GPU box); skipped elsewhere - `tests/test_oracle_golden.py` pins the same oracle to
the committed outputs of the reference and always runs."""
import numpy as np
import pytest
import torch

from oracle.iql_oracle import IQLOracle, OracleHParams, nets_from_reference_modules
from oracle.ref_import import load_reference_iql, reference_available

pytestmark = pytest.mark.skipif(not reference_available(), reason="the reference is not importable here")

HP = dict(discount=0.97, tau=0.01, beta=5.0, iql_tau=0.8, vf_lr=1e-3, qf_lr=5e-4, actor_lr=2e-4, max_steps=7)


@pytest.mark.parametrize("mode", ["fp32", "autocast"])
@pytest.mark.parametrize("det", [False, True])
def test_oracle_tracks_the_imported_reference(mode, det):
    ref = load_reference_iql()
    S, A, H, B, K = 13, 5, 48, 96, 9  # K > max_steps: the cosine schedule passes its restart branch
    torch.manual_seed(21)
    qf, vf = ref.TwinQ(S, A, H), ref.ValueFunction(S, H)
    actor = (ref.DeterministicPolicy if det else ref.GaussianPolicy)(S, A, 1.0, H)
    orc = IQLOracle(*nets_from_reference_modules(qf, vf, actor), OracleHParams(**HP, deterministic=det), mode=mode)
    mk = lambda m, lr: torch.optim.Adam(m.parameters(), lr=lr)
    a_opt = mk(actor, HP["actor_lr"])
    tr = ref.ImplicitQLearning(1.0, actor, a_opt, qf, mk(qf, HP["qf_lr"]), vf, mk(vf, HP["vf_lr"]),
                               iql_tau=HP["iql_tau"], beta=HP["beta"], max_steps=HP["max_steps"],
                               discount=HP["discount"], tau=HP["tau"], device="cpu")
    rng = np.random.default_rng(8)
    for k in range(K):
        obs = torch.tensor(rng.standard_normal((B, S)), dtype=torch.float32)
        act = torch.tensor(rng.uniform(-1, 1, (B, A)), dtype=torch.float32)
        rew = torch.tensor(rng.standard_normal(B), dtype=torch.float32)
        nobs = torch.tensor(rng.standard_normal((B, S)), dtype=torch.float32)
        done = torch.tensor((rng.uniform(size=B) < 0.2).astype(np.float32))
        got = orc.step(obs, act, rew, nobs, done)
        if mode == "autocast":
            want = tr.train([obs, act, rew[:, None], nobs, done[:, None]])
        else:
            tr.total_it += 1
            want = {}
            with torch.no_grad():
                next_v = tr.vf(nobs)
            adv = tr._update_v(obs, act, want)
            tr._update_q(next_v, obs, act, rew, done, want)
            tr._update_policy(adv, obs, act, want)
        for key in ("value_loss", "q_loss", "actor_loss"):
            assert got[key] == pytest.approx(want[key], rel=1e-6, abs=1e-8), (k, key)
        assert orc.a_opt.lr == pytest.approx(a_opt.param_groups[0]["lr"], rel=1e-12)
    e = orc.export()
    for net, mod in (("q1", qf.q1), ("q2", qf.q2), ("v", vf.v), ("actor", actor.net), ("qt1", tr.q_target.q1),
                     ("qt2", tr.q_target.q2)):
        lins = [m for m in mod.net if isinstance(m, torch.nn.Linear)]
        for i, lin in enumerate(lins):
            np.testing.assert_allclose(e[f"{net}.w{i}"], lin.weight.detach().numpy(), rtol=1e-5, atol=1e-7)
            np.testing.assert_allclose(e[f"{net}.b{i}"], lin.bias.detach().numpy(), rtol=1e-5, atol=1e-7)
    if not det:
        np.testing.assert_allclose(e["log_std"], actor.log_std.detach().numpy(), rtol=1e-5, atol=1e-7)
