"""GPU parity of the preference-transformer training step: `iqlpref_b200.pt_training.PTTrainer`
(`iqlb_pt_train_step` / `iqlb_pt_eval`, csrc/pt_train.cu) against

* fixtures written by the unmodified reference `PTTrainer` on its own `PT` with every dropout at 0
  (gp_reward-priors/optbnn/training/training.py:308-416; tests/golden/pt_train.npz, oracle/gen_golden_pt.py): first-step
  gradients of every parameter, per-step loss / accuracy, `evaluation()`, weights and Adam moments after K steps;
* the torch oracle (oracle/pt_train_oracle.py, pinned to the same fixtures) on fresh seeded batches at larger shapes, in
  both arithmetic modes, and WITH dropout: the oracle takes the documented Philox keep masks (include/iqlb.h) as explicit
  inputs, the CUDA step regenerates them from (seed, site, Adam step).

Tolerances: loss 1e-4 relative, accuracy exact, gradients 1e-3 of the tensor's largest entry with a floor at 1e-3 of the
whole gradient's largest entry, parameters after K Adam steps as in tests/test_pt_oracle_golden.py (3e-5 absolute except
rounding-noise gradients, bounded by 2 K lr).  `precision="split_bf16"`: first-step loss 2e-4, gradients 1.5e-2 of a tensor's
largest entry (measured 6e-3; the bar of the MR trainer's tensor-core mode, DESIGN.md §2: a Bradley-Terry gradient near the
ln 2 plateau is a cancelling sum over thousands of rows that amplifies the 2^-17 product error), losses after Adam steps at
lr 1e-3 2e-3 (Adam turns the rounding of near-zero gradients into +-lr moves)."""
import os

import numpy as np
import pytest
import torch

from oracle.gen_golden_pt import CASES, load_case
from oracle.pt_train_oracle import PTTrainerOracle, dropout_masks, synth_pt_batch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def golden(golden_dir):
    return np.load(os.path.join(golden_dir, "pt_train.npz"))


def build(w0, S, A, n_ts, E, P, heads, I, L, eps, pool, lr=3e-4, precision="fp32", drops=(0.0, 0.0, 0.0), seed=7):
    from iqlpref_b200.pt_training import PTTrainer
    from iqlpref_b200.reward_models.pref_transformer import PT
    net = PT(state_dim=S, action_dim=A, max_episode_steps=n_ts - 1, embd_dim=E, pref_attn_embd_dim=P, num_heads=heads,
             attn_dropout=drops[1], resid_dropout=drops[2], intermediate_dim=I, num_layers=L, embd_dropout=drops[0],
             max_pos=64, eps=eps)
    net.resid_dropout_p = drops[2]  # the fixtures force every nn.Dropout to 0 (gen_golden_pt.build)
    sd = net.state_dict()
    with torch.no_grad():
        for k, v in w0.items():
            sd[k].copy_(torch.from_numpy(np.asarray(v)))
    opt = torch.optim.Adam(net.parameters(), lr=lr)
    return PTTrainer(net, opt, device="cuda", bt_pool=pool, precision=precision, seed=seed), net, opt


def tb(batch):
    return [torch.from_numpy(np.asarray(x)) for x in batch]


def assert_grads_close(got, want, rel=1e-3, outliers=0.0):
    """`outliers`: fraction of a tensor's elements (at least one) that may miss the bar by up to 3x (split-bf16: a hidden unit
    whose pre-activation is within rounding of 0 switches its ReLU between the two arithmetics)."""
    floor = max(float(np.abs(np.asarray(g)).max()) for g in want.values())
    assert set(got) == set(want)
    for k, w in want.items():
        w = np.asarray(w)
        g = got[k].detach().cpu().numpy()
        atol = rel * max(float(np.abs(w).max()), 1e-3 * floor) + 1e-9
        d = np.abs(g - w)
        assert d.max() <= (3 * atol if outliers else atol), (k, float(d.max()), atol)
        assert np.sum(d > atol) <= (max(1, outliers * d.size) if outliers else 0), (k, int(np.sum(d > atol)))


def assert_weights_close(got, want, g0, K, lr=3e-4):
    gmax = max(float(np.abs(g).max()) for g in g0.values())
    for k, w in want.items():
        d = np.abs(got[k].detach().cpu().numpy() - w)
        assert d.max() <= 2 * K * lr, k
        real = np.abs(g0[k]) > 1e-5 * gmax
        assert np.sum(d[real] > 3e-5) <= max(1, 0.02 * d.size), k


@pytest.mark.parametrize("case", sorted(CASES))
def test_pt_step_matches_the_reference_fixtures(golden, case):
    c = load_case(golden, case)
    n_ts = c["w0"]["timestep_embed.weight"].shape[0]
    tr, net, opt = build(c["w0"], c["S"], c["A"], n_ts, c["E"], c["P"], c["heads"], c["I"], c["L"], c["eps"], c["pool"])
    loss, acc, grad = tr.loss_and_grad(tb(c["batches"][0]))
    assert loss == pytest.approx(c["loss"][0], rel=1e-4) and acc == pytest.approx(c["acc"][0], abs=1e-7)
    assert_grads_close(tr.grad_dict(grad), c["g0"])
    for k in range(c["K"]):
        log = tr.train(tb(c["batches"][k]))
        assert log["training_loss"] == pytest.approx(c["loss"][k], rel=1e-4), k
        assert log["training_acc"] == pytest.approx(c["acc"][k], abs=1e-7), k
    ev = tr.evaluation(tb(c["batches"][c["K"]]))
    assert ev["eval_loss"] == pytest.approx(c["eval"][0], rel=1e-4) and ev["eval_acc"] == pytest.approx(c["eval"][1], abs=1e-7)
    sd = {k: v for k, v in net.state_dict().items() if "causal_bias" not in k}
    assert_weights_close(sd, c["wK"], c["g0"], c["K"])
    # Adam moments in the reference's optimizer.state_dict() format (index = position in net.parameters())
    st = tr.state_dict()["optimizer"]["state"]
    names = [n for n, _ in net.named_parameters()]
    for i, n in enumerate(names):
        assert float(st[i]["step"]) == c["K"]
        m = st[i]["exp_avg"].cpu().numpy()
        scale = max(float(np.abs(c["mK"][n]).max()), 1e-8)
        np.testing.assert_allclose(m, c["mK"][n], rtol=0, atol=2e-3 * scale + 1e-9, err_msg=n)


SHAPES = {
    # name: (S, A, T, B, E, P, heads, I, L, pool)
    "l2_ragged": (11, 4, 25, 6, 64, 32, 2, 128, 2, "mean"),
    "l1_wide": (29, 8, 20, 8, 128, 128, 2, 256, 1, "sum"),
    # 2T = 128 tokens, head_dim 64, T = 64 >= the tensor-core threshold of the grouped attention GEMMs (pt_train.cu ptt_batched):
    # with precision="split_bf16" S = Q K^T, O = P V and their four gradients run on tcgen05, per (sequence, head)
    "l1_tc_attn": (11, 4, 64, 4, 128, 64, 2, 256, 1, "mean"),
}


def random_weights(rng, S, A, n_ts, E, P, I, L):
    from iqlpref_b200.pt_training import pt_parameter_names
    shapes = {"state_linear.weight": (E, S), "state_linear.bias": (E,), "action_linear.weight": (E, A), "action_linear.bias": (E,),
              "timestep_embed.weight": (n_ts, E), "pref_linear.weight": (2 * P + 1, E), "pref_linear.bias": (2 * P + 1,)}
    for l in range(L):
        p = f"gpt.layers.{l}."
        shapes.update({p + "attention.in_linear.weight": (3 * E, E), p + "attention.in_linear.bias": (3 * E,),
                       p + "attention.out_linear.weight": (E, E), p + "attention.out_linear.bias": (E,),
                       p + "mlp.in_linear.weight": (I, E), p + "mlp.in_linear.bias": (I,),
                       p + "mlp.out_linear.weight": (E, I), p + "mlp.out_linear.bias": (E,)})
    w = {}
    for name in pt_parameter_names(L):
        if "layer_norm" in name:
            w[name] = (1.0 + 0.2 * rng.standard_normal(E) if name.endswith("weight") else 0.2 * rng.standard_normal(E)).astype(np.float32)
        elif name == "timestep_embed.weight":
            w[name] = rng.standard_normal(shapes[name]).astype(np.float32)
        elif name.endswith("bias"):
            w[name] = (0.2 * rng.standard_normal(shapes[name])).astype(np.float32)
        else:
            fan_in = shapes[name][1]
            w[name] = (rng.standard_normal(shapes[name]) / np.sqrt(fan_in)).astype(np.float32)
    return w


@pytest.mark.parametrize("precision,rel_g,rel_l", [("fp32", 1e-3, 1e-4), ("split_bf16", 1.5e-2, 2e-4)])
@pytest.mark.parametrize("drops", [(0.0, 0.0, 0.0), (0.1, 0.1, 0.1)], ids=["nodrop", "dropout"])
@pytest.mark.parametrize("shape", sorted(SHAPES))
def test_pt_step_matches_oracle_with_injected_masks(shape, drops, precision, rel_g, rel_l):
    S, A, T, B, E, P, heads, I, L, pool = SHAPES[shape]
    rng = np.random.default_rng(11 + sorted(SHAPES).index(shape))
    n_ts, eps, seed, lr = T + 3, 0.1, 99, 1e-3
    w0 = random_weights(rng, S, A, n_ts, E, P, I, L)
    tr, net, opt = build(w0, S, A, n_ts, E, P, heads, I, L, eps, pool, lr=lr, precision=precision, drops=drops, seed=seed)
    orc = PTTrainerOracle(w0, heads, eps, lr=lr, bt_pool=pool)
    K = 3
    for k in range(K):
        batch = synth_pt_batch(rng, B, T, S, A)
        masks = dropout_masks(seed, k + 1, 2 * B, T, E, heads, L, drops[0], drops[1], drops[2], drops[2])
        if k == 0:
            want_loss, want_acc, want_g = orc.loss_and_grads(batch, masks)
            loss, acc, grad = tr.loss_and_grad(tb(batch), training=True)
            assert loss == pytest.approx(want_loss, rel=rel_l)
            assert acc == pytest.approx(want_acc, abs=1e-7)
            assert_grads_close(tr.grad_dict(grad), {k_: v.numpy() for k_, v in want_g.items()}, rel=rel_g,
                               outliers=0.0 if precision == "fp32" else 1e-3)
            g_first = want_g
        want = orc.train(batch, masks)
        got = tr.train(tb(batch))
        step_rel = rel_l if (precision == "fp32" or k == 0) else 2e-3
        assert got["training_loss"] == pytest.approx(want["training_loss"], rel=step_rel), k
        assert got["training_acc"] == pytest.approx(want["training_acc"], abs=1e-7), k
    batch = synth_pt_batch(rng, B, T, S, A)
    ev, want_ev = tr.evaluation(tb(batch)), orc.evaluation(batch)
    assert ev["eval_loss"] == pytest.approx(want_ev["eval_loss"], rel=5e-4 if precision == "fp32" else 2e-3)
    got_w = {k: v for k, v in net.state_dict().items() if "causal_bias" not in k}
    # elements whose gradient is rounding noise around an exact 0 (key biases of both attentions: a softmax ignores a
    # constant added to all its scores) move by ~lr per step in either direction under Adam; recognised by the oracle's
    # first-step gradient, bounded by 2 K lr (tests/test_pt_oracle_golden.py)
    gmax = max(float(v.abs().max()) for v in g_first.values())
    for k, w in orc.weights().items():
        d = np.abs(got_w[k].cpu().numpy() - w)
        assert d.max() <= 2 * K * lr, k
        real = np.abs(g_first[k].numpy()) > 1e-5 * gmax
        assert np.sum(d[real] > (1e-4 if precision == "fp32" else 4e-4)) <= max(1, 0.03 * d.size), k


def test_pt_trainer_checkpoint_round_trip_and_relabel(golden, tmp_path):
    """`state_dict()` in the reference's format, `fit_pt` -> best_model.pt -> `load_pt_reward_model` ->
    `qlearning_dataset_pt` (BASELINE configs[2]: train the PT, relabel, hand the rewards to IQL)."""
    from iqlpref_b200 import relabel
    from iqlpref_b200.pt_training import fit_pt
    c = load_case(golden, "l1_mean")
    n_ts = c["w0"]["timestep_embed.weight"].shape[0]
    args = (c["S"], c["A"], n_ts, c["E"], c["P"], c["heads"], c["I"], c["L"], c["eps"], c["pool"])
    tr, net, opt = build(c["w0"], *args)
    batches = [tb(b) for b in c["batches"]]
    hist = fit_pt(tr, batches[:2], batches[2:3], epochs=2, eval_every=1, checkpoints_path=str(tmp_path), criteria_key="loss")
    assert len(hist) == 3 and os.path.isfile(tmp_path / "best_model.pt")
    assert not os.path.isfile(tmp_path / "checkpoint_0.pt")  # the PT script writes no per-epoch snapshots
    sd = tr.state_dict()
    assert set(sd) == {"net", "optimizer"} and len(sd["optimizer"]["state"]) == len(list(net.parameters()))
    tr2, net2, _ = build(c["w0"], *args)
    tr2.load_state_dict(sd)
    assert tr2.step_count == tr.step_count and torch.equal(tr2.flat, tr.flat) and torch.equal(tr2.exp_avg, tr.exp_avg)
    a, b = tr.train(batches[0]), tr2.train(batches[0])
    assert a == b and torch.equal(tr2.flat, tr.flat)
    # the trained network relabels a dataset: its parameters are views of the trainer's buffer (no copy), and the
    # packed-weight cache of the relabel path follows the step; best_model.pt loads through load_pt_reward_model
    import yaml
    from iqlpref_b200.iql import load_pt_reward_model, qlearning_dataset_pt

    class Env:
        _max_episode_steps = 15

    rng = np.random.default_rng(5)
    n = 40
    ds = {"observations": rng.standard_normal((n, c["S"])).astype(np.float32),
          "actions": rng.uniform(-1, 1, (n, c["A"])).astype(np.float32),
          "next_observations": rng.standard_normal((n, c["S"])).astype(np.float32),
          "rewards": np.zeros(n, np.float32), "terminals": np.zeros(n, np.float32)}
    ds["terminals"][[13, 29]] = 1
    net.eval()
    before = qlearning_dataset_pt(Env(), net, query_length=c["T"], dataset=ds)["rewards"].copy()
    (tmp_path / "config.yaml").write_text(yaml.safe_dump({"num_heads": c["heads"], "intermediate_dim": c["I"],
                                                          "model_eps": c["eps"]}))
    torch.save(tr.state_dict(), tmp_path / "best_model.pt")
    loaded = load_pt_reward_model(str(tmp_path), "cuda")
    np.testing.assert_allclose(qlearning_dataset_pt(Env(), loaded, query_length=c["T"], dataset=ds)["rewards"], before,
                               rtol=1e-6, atol=1e-7)
    tr.train(batches[1])
    net.eval()
    after = qlearning_dataset_pt(Env(), net, query_length=c["T"], dataset=ds)["rewards"]
    assert np.abs(after - before).max() > 0


def test_pt_trainer_rejects_bad_input(golden):
    c = load_case(golden, "l1_mean")
    n_ts = c["w0"]["timestep_embed.weight"].shape[0]
    tr, net, opt = build(c["w0"], c["S"], c["A"], n_ts, c["E"], c["P"], c["heads"], c["I"], c["L"], c["eps"], c["pool"])
    batch = tb(c["batches"][0])
    bad = list(batch)
    bad[2] = batch[2].clone()
    bad[2][0, -1] = n_ts
    with pytest.raises(IndexError):
        tr.train(bad)
    bad = list(batch)
    bad[0] = batch[0][:, :, :-1]
    with pytest.raises((RuntimeError, ValueError)):
        tr.train(bad)
    from iqlpref_b200.pt_training import PTTrainer
    with pytest.raises(RuntimeError):
        PTTrainer(net, opt, device="cpu")
    with pytest.raises(NotImplementedError):
        PTTrainer(net, torch.optim.SGD(net.parameters(), lr=0.1), device="cuda")
