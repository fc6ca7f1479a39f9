"""GPU parity of the bf16 tcgen05 step.  bf16 operands (weights, activations, output gradients rounded
to bf16; fp32 accumulation in TMEM; fp32 parameters, Adam state and target) - the arithmetic class of
the reference's as-shipped autocast path (iql.py:650).

The parity target for this dtype is the reference ITSELF run under CUDA autocast on a B200 (O-ac-cuda,
SURVEY.md 8c): `tests/golden/iql_cfg{1,2,4det}_autocast_cuda.npz` were written on the GPU box by
`scripts/gen_cuda_golden.py` from the unmodified reference (`baseline/_ref`, oracle/make_ref.py), next to its
fp32 run on the same device (`*_fp32_cuda.npz`).  The tolerance is DERIVED FROM THOSE TWO FILES, not assumed:
the reference's own autocast run deviates from its own fp32 run by 0.7e-2..1.1e-2 on the value loss, 1e-3..2e-3
on the q loss, 2e-3..1e-2 on the actor loss and 2e-2..1.8e-1 of a tensor's largest entry on the parameters
after 20 steps (profiles/r02_cuda_autocast_spread.json); precision="bf16" must agree with the reference's
autocast output to within 1.5x that spread, per loss column and on the worst parameter tensor (observed: 0.55x
to 1.12x - it is as close to the as-shipped reference as the reference's fp32 statement is, and closer to the
fp32 statement than the as-shipped reference is).  When `baseline/_ref` travelled to the box the reference is
also run live and must reproduce its fixture.

Secondary checks against this repo's own fp32 path (same step, same inputs):
  * per-tensor gradients: relative L2 error <= 8e-2, losses <= 3e-2 relative (+1e-4 abs) over K steps.
The 1e-3 bar of BASELINE.json is met by precision="fp32" (tests/test_gpu_iql_step.py)."""
import os

import numpy as np
import pytest
import torch
import yaml

from oracle.iql_oracle import IQLOracle, OracleHParams, nets_from_reference_modules

pytestmark = pytest.mark.gpu

HP = dict(discount=0.99, tau=0.005, beta=3.0, iql_tau=0.7, vf_lr=3e-4, qf_lr=3e-4, actor_lr=3e-4, max_steps=1000)


def _batch(rng, B, S, A):
    return (torch.tensor(rng.standard_normal((B, S)), dtype=torch.float32),
            torch.tensor(rng.uniform(-1, 1, (B, A)), dtype=torch.float32),
            torch.tensor(rng.standard_normal(B) * 0.5 + 0.5, dtype=torch.float32),
            torch.tensor(rng.standard_normal((B, S)), dtype=torch.float32),
            torch.tensor((rng.uniform(size=B) < 0.1).astype(np.float32)))


def _engines(biql, S, A, det, n_agents, B=256):
    engs = {}
    for prec in ("fp32", "bf16"):
        engs[prec] = biql.BatchedIQL(S, A, n_agents, batch_size=B, deterministic=det, hparams=[HP] * n_agents,
                                     precision=prec)
    mods = []
    for a in range(n_agents):
        torch.manual_seed(10 + a)
        qf, vf = biql.TwinQ(S, A), biql.ValueFunction(S)
        actor = (biql.DeterministicPolicy if det else biql.GaussianPolicy)(S, A, 1.0)
        mods.append((qf, vf, actor))
        for e in engs.values():
            e.adopt(a, qf, vf, actor, bind=False)
    return engs, mods


# (111, 8): the Ant shape - every stream reads two input blocks (45, 24: only the Q streams do)
@pytest.mark.parametrize("S,A,det", [(17, 6, False), (29, 8, False), (45, 24, True), (111, 8, False)])
def test_bf16_forward_and_gradients_match_fp32_path(S, A, det):
    from iqlpref_b200 import iql as biql
    N = 2
    engs, _ = _engines(biql, S, A, det, N)
    rng = np.random.default_rng(S)
    rows = torch.stack([biql.pack_rows(*_batch(rng, 256, S, A), engs["fp32"].R) for _ in range(N)]).cuda()
    for e in engs.values():
        e.train(rows)
    torch.cuda.synchronize()
    f, b = engs["fp32"], engs["bf16"]
    ho_f = f.debug_tensor(2).view(N, 6 + A, 256).cpu()
    ho_b = b.debug_tensor(2).view(N, 6 + A, 256).cpu()
    assert torch.isfinite(ho_b).all()
    for r in range(6 + A):
        ref = ho_f[:, r]
        err = (ho_b[:, r] - ref).norm() / ref.norm().clamp_min(1e-6)
        assert err < 2e-2, f"head_out row {r}: rel L2 err {err:.3e}"
    g_f = f.debug_tensor(6).view(N, -1).cpu()
    g_b = b.debug_tensor(6).view(N, -1).cpu()
    assert torch.isfinite(g_b).all()
    for key, (off, shape) in f._layout.items():
        n = int(np.prod(shape))
        for a in range(N):
            ref, got = g_f[a, off:off + n], g_b[a, off:off + n]
            err = (got - ref).norm() / ref.norm().clamp_min(1e-12)
            assert err < 8e-2, f"agent {a} grad {key}: rel L2 err {err:.3e} (|ref| {ref.norm():.3e})"
    np.testing.assert_allclose(b.losses.cpu().numpy(), f.losses.cpu().numpy(), rtol=2e-2, atol=1e-4)


def test_bf16_k_steps_track_the_fp32_oracle():
    from iqlpref_b200 import iql as biql
    S, A, N, K = 29, 8, 2, 12
    engs, mods = _engines(biql, S, A, False, N)
    eng = engs["bf16"]
    oracles = [IQLOracle(*nets_from_reference_modules(*m), OracleHParams(**HP), mode="fp32") for m in mods]
    rng = np.random.default_rng(3)
    for k in range(K):
        rows, want = [], []
        for a in range(N):
            bt = _batch(rng, 256, S, A)
            w = oracles[a].step(*bt)
            want.append([w["value_loss"], w["q_loss"], w["actor_loss"]])
            rows.append(biql.pack_rows(*bt, eng.R))
        eng.train(torch.stack(rows).cuda())
        got = eng.losses.cpu().numpy()
        np.testing.assert_allclose(got, np.array(want), rtol=3e-2, atol=1e-4, err_msg=f"step {k}")
    for a in range(N):
        e, o = eng.export(a), oracles[a].export()
        for key in ("q1.w1", "v.w0", "actor.w2", "qt2.w1", "q2.b1", "actor.b0"):
            scale = np.abs(o[key]).max()
            assert np.abs(e[key] - o[key]).max() <= 2e-2 * scale + K * 3e-4, key


def test_bf16_sampled_step_and_shadow_refresh():
    """On-device sampling feeds the same kernels; the weight images follow Adam (no drift vs a rebuild)."""
    from iqlpref_b200 import iql as biql
    S, A, N = 29, 8, 3
    engs, _ = _engines(biql, S, A, False, N)
    eng = engs["bf16"]
    buf = biql.ReplayBuffer(S, A, 5000)
    g = torch.Generator(device="cuda").manual_seed(0)
    buf._table[:, :2 * S + A + 2] = torch.randn((5000, 2 * S + A + 2), generator=g, device="cuda")
    buf._table[:, 2 * S + A + 1] = 0
    buf._size = buf._pointer = 5000
    eng.rng[:, 0] = torch.arange(N, device="cuda") + 5
    eng.attach_replay(buf)
    for _ in range(5):
        eng.train(None)
    a = eng.losses.clone()
    assert torch.isfinite(a).all()
    # rebuilding the images from the fp32 parameters must not change the next step
    state = [t.clone() for t in (eng.params, eng.exp_avg, eng.exp_avg_sq, eng.target, eng.it, eng.actor_lr, eng.rng)]
    eng.train(None)
    l1 = eng.losses.clone()
    for t, s in zip((eng.params, eng.exp_avg, eng.exp_avg_sq, eng.target, eng.it, eng.actor_lr, eng.rng), state):
        t.copy_(s)
    eng._sync_shadow()
    eng.train(None)
    assert torch.equal(l1, eng.losses)


def test_bf16_lookahead_sampling_changes_nothing(monkeypatch):
    """The gather of step t+1 running under the optimizer kernel of step t on a side stream (IQLB_PRESAMPLE=1) must leave
    losses, parameters and RNG offsets bit-identical to the in-order schedule (0, the default), also across an explicit-batch
    step and a table swap."""
    from iqlpref_b200 import iql as biql
    S, A, N = 29, 8, 4
    buf = biql.ReplayBuffer(S, A, 7000)
    g = torch.Generator(device="cuda").manual_seed(1)
    buf._table[:, :2 * S + A + 2] = torch.randn((7000, 2 * S + A + 2), generator=g, device="cuda")
    buf._table[:, 2 * S + A + 1] = 0
    buf._size = buf._pointer = 7000
    buf2 = biql.ReplayBuffer(S, A, 3000)
    buf2._table[:, :2 * S + A + 2] = torch.randn((3000, 2 * S + A + 2), generator=g, device="cuda")
    buf2._table[:, 2 * S + A + 1] = 0
    buf2._size = buf2._pointer = 3000
    rng = np.random.default_rng(2)
    rows = torch.stack([biql.pack_rows(*_batch(rng, 256, S, A), buf._stride) for _ in range(N)]).cuda()
    out = {}
    for mode in ("0", "1"):
        monkeypatch.setenv("IQLB_PRESAMPLE", mode)
        eng = biql.BatchedIQL(S, A, N, batch_size=256, hparams=[HP] * N, precision="bf16", seeds=[7, 8, 9, 10])
        for a in range(N):
            torch.manual_seed(20 + a)
            eng.adopt(a, biql.TwinQ(S, A), biql.ValueFunction(S), biql.GaussianPolicy(S, A, 1.0), bind=False)
        eng.attach_replay(buf)
        trace = []
        for k in range(9):
            if k == 4:
                eng.train(rows)          # explicit batch between sampled steps
            else:
                if k == 6:
                    eng.attach_replay(buf2)  # table swap with a look-ahead batch pending
                eng.train(None)
            trace.append(eng.losses.clone())
        torch.cuda.synchronize()
        out[mode] = (torch.stack(trace), eng.params.clone(), eng.rng.clone())
    for x, y in zip(out["0"], out["1"]):
        assert torch.equal(x, y)
    assert out["1"][2][:, 1].tolist() == [32] * N  # eight sampled steps, four Philox draws each


# ------------------------------------------------------------------------------------------------
# the benchmarked dtype against the reference's own CUDA-autocast output
# ------------------------------------------------------------------------------------------------
CFGS = {"cfg1": (17, 6, False), "cfg2": (29, 8, False), "cfg4det": (45, 24, True)}
SPREAD_FACTOR = 1.5  # allowed multiple of the reference's own autocast <-> fp32 deviation (module docstring)


def _loss_dev(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return (np.abs(a - b) / np.maximum(np.abs(b), 1e-6)).max(axis=0)


def _param_dev(fa, fb):
    """worst tensor: max |a - b| over the fixtures' strided samples, normalised by the tensor's largest entry"""
    worst = 0.0
    for k in fb:
        a, b = np.asarray(fa[k], np.float64), np.asarray(fb[k], np.float64)
        worst = max(worst, float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-12)))
    return worst


def _fixture(golden_dir, cfg, mode):
    d = np.load(os.path.join(golden_dir, f"iql_{cfg}_{mode}_cuda.npz"))
    final = {k[len("final."):-len("#sample")]: d[k] for k in d.files if k.startswith("final.") and k.endswith("#sample")}
    return d, final


def _run_bf16(biql, cfg, d):
    """K steps of precision="bf16" on the fixture's seed, table and index stream (oracle/gen_golden.py draws)."""
    S, A, det = CFGS[cfg]
    hp = yaml.safe_load(str(d["hp"]))
    torch.manual_seed(int(d["seed"]))
    qf, vf = biql.TwinQ(S, A), biql.ValueFunction(S)
    actor = (biql.DeterministicPolicy if det else biql.GaussianPolicy)(S, A, 1.0)
    for nm, mod in (("qf", qf), ("vf", vf), ("actor", actor)):  # same initialisation as the reference's run
        for key, val in mod.state_dict().items():
            np.testing.assert_array_equal(val.numpy().reshape(-1)[::97], d[f"init.{nm}.{key}#sample"])
    eng = biql.BatchedIQL(S, A, 1, batch_size=256, deterministic=det, hparams=[hp], precision="bf16")
    eng.adopt(0, qf, vf, actor)
    n_rows = int(d["n_rows"])
    rng = np.random.default_rng(int(d["seed"]) + 1000)
    tab = {"obs": rng.standard_normal((n_rows, S)).astype(np.float32),
           "act": rng.uniform(-1, 1, (n_rows, A)).astype(np.float32),
           "rew": (rng.standard_normal(n_rows) * 0.5 + 0.5).astype(np.float32),
           "next_obs": rng.standard_normal((n_rows, S)).astype(np.float32),
           "done": (rng.uniform(size=n_rows) < 0.05).astype(np.float32)}
    K = d["idx"].shape[0]
    losses = np.zeros((K, 3))
    for k in range(K):
        i = torch.from_numpy(d["idx"][k])
        rows = biql.pack_rows(torch.from_numpy(tab["obs"])[i], torch.from_numpy(tab["act"])[i],
                              torch.from_numpy(tab["rew"])[i], torch.from_numpy(tab["next_obs"])[i],
                              torch.from_numpy(tab["done"])[i], eng.R)
        eng.train(rows.cuda().unsqueeze(0))
        losses[k] = eng.losses[0].cpu().numpy()
    final = {}
    for nm, mod in (("qf", qf), ("vf", vf), ("actor", actor)):
        for key, val in mod.state_dict().items():
            final[f"{nm}.{key}"] = val.detach().cpu().numpy().reshape(-1)[::97]
    e = eng.export(0)
    for net in ("q1", "q2"):
        for i in range(3):
            final[f"q_target.{net}.net.{2 * i}.weight"] = e[f"qt{net[1]}.w{i}"].reshape(-1)[::97]
            final[f"q_target.{net}.net.{2 * i}.bias"] = e[f"qt{net[1]}.b{i}"].reshape(-1)[::97]
    assert abs(float(eng.actor_lr[0].item()) - float(d["lr_final"])) < 1e-12
    return losses, final


@pytest.mark.parametrize("cfg", list(CFGS))
def test_bf16_step_matches_reference_cuda_autocast(golden_dir, cfg):
    """precision="bf16" against what the unmodified reference produces under CUDA autocast on a B200 (K = 20 steps,
    losses per step and parameters + target after the last one), at 1.5x the reference's own autocast <-> fp32 spread."""
    from iqlpref_b200 import iql as biql
    d_ac, f_ac = _fixture(golden_dir, cfg, "autocast")
    d_32, f_32 = _fixture(golden_dir, cfg, "fp32")
    spread_loss, spread_param = _loss_dev(d_ac["losses"], d_32["losses"]), _param_dev(f_ac, f_32)
    assert (spread_loss > 5e-4).all() and spread_param > 1e-3, "the two reference runs are suspiciously identical"
    losses, final = _run_bf16(biql, cfg, d_ac)
    assert set(final) == set(f_ac)
    dev_loss, dev_param = _loss_dev(losses, d_ac["losses"]), _param_dev(final, f_ac)
    print(f"{cfg}: losses {dev_loss} vs reference spread {spread_loss}; params {dev_param:.3e} vs {spread_param:.3e}")
    assert (dev_loss <= SPREAD_FACTOR * spread_loss).all(), (dev_loss, spread_loss)
    assert dev_param <= SPREAD_FACTOR * spread_param, (dev_param, spread_param)
    # ... and it is at least as close to the reference's fp32 statement as the as-shipped reference is
    assert (_loss_dev(losses, d_32["losses"]) <= SPREAD_FACTOR * spread_loss).all()
    assert _param_dev(final, f_32) <= SPREAD_FACTOR * spread_param


@pytest.mark.parametrize("cfg", list(CFGS))
def test_reference_cuda_autocast_live_reproduces_its_fixture(golden_dir, cfg):
    """With `baseline/_ref` on the box (oracle/make_ref.py): run the unmodified reference under CUDA autocast now and
    check the committed fixture is what it produces (losses to 1e-5: cuBLAS may pick another kernel on another driver)."""
    from oracle.ref_import import load_reference_iql, reference_available
    if not reference_available():
        pytest.skip("baseline/_ref did not travel to this box: the committed CUDA fixtures stand in")
    from oracle.gen_golden import run_reference_steps
    ref = load_reference_iql()
    d, f_fix = _fixture(golden_dir, cfg, "autocast")
    S, A, det = CFGS[cfg]
    hp = yaml.safe_load(str(d["hp"]))
    _, idx, _, final, losses, _ = run_reference_steps(ref, S, A, 256, 256, d["idx"].shape[0], int(d["n_rows"]),
                                                      int(d["seed"]), "autocast", det, hp, device="cuda")
    np.testing.assert_array_equal(idx, d["idx"])
    np.testing.assert_allclose(losses, d["losses"], rtol=1e-5, atol=1e-8)
    live = {k: v.reshape(-1)[::97] for k, v in final.items()}
    assert _param_dev(live, f_fix) < 1e-4
