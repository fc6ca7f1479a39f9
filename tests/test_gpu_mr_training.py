"""GPU parity of the MR preference-model training step: `iqlpref_b200.mr_training.MRTrainer`
(`iqlb_mr_train_step` / `iqlb_mr_eval`) against fixtures written by the unmodified reference `MRTrainer`
(gp_reward-priors/optbnn/training/training.py:27-114; tests/golden/mr_train.npz, oracle/gen_golden_mr.py) and
against the numpy oracle on fresh seeded batches for every activation, both pooling modes, ragged / empty masks,
NaN rewards and the production batch shape.  Tolerances: loss 1e-4 relative, gradients 1e-3 relative to the tensor's
largest entry (floored at 1e-2 of the whole gradient's largest entry), accuracy exact, parameters after K Adam steps 3e-5 absolute except elements whose gradient is rounding
noise (Adam moves those by ~lr per step in either direction; at most 2 %, bounded by 2*K*lr)."""
import os

import numpy as np
import pytest
import torch

from oracle.gen_golden_mr import CASES, load_case
from oracle.mr_train_oracle import MRTrainerOracle, mr_loss_and_grads, synth_batch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def golden(golden_dir):
    return np.load(os.path.join(golden_dir, "mr_train.npz"))


def make_trainer(weights, act, pool, lr=3e-4, precision="fp32"):
    from iqlpref_b200.mr_training import MRTrainer
    from iqlpref_b200.reward_models.q_mlp import MLP
    d_in = weights[0].shape[0]
    widths = [w.shape[1] for w in weights[0:-2:2]]
    net = MLP(d_in, 1, widths, act)
    with torch.no_grad():
        for p, w in zip(net.parameters(), weights):
            p.copy_(torch.from_numpy(np.asarray(w)).reshape(p.shape))
    opt = torch.optim.Adam(net.parameters(), lr=lr)
    return MRTrainer(net, opt, num_datapoints=1, device="cuda", bt_pool=pool, precision=precision), net, opt


def tb(batch):
    return [torch.from_numpy(np.asarray(x)) for x in batch]


def split(flat, weights):
    out, off = [], 0
    flat = flat.detach().cpu().numpy()
    for w in weights:
        out.append(flat[off:off + w.size].reshape(w.shape)); off += w.size
    return out


def assert_grads_close(got, want, rel=1e-3):
    # per tensor, relative to its largest entry - with a floor at 1e-2 of the whole gradient's largest entry: the
    # output bias has a Bradley-Terry gradient of exactly 0 in real arithmetic, so both sides hold rounding noise there
    floor = 1e-2 * max(float(np.abs(np.asarray(gr)).max()) for gr in want)
    for i, (g, gr) in enumerate(zip(got, want)):
        gr = np.asarray(gr).reshape(g.shape)
        if i == len(want) - 1:  # the output bias: with mean pooling its gradient is rounding noise around an exact 0
            # (a cancelling sum of 2*B*T terms whose magnitudes add up to <= 1: a few fp32 ulps of 1 in absolute terms)
            np.testing.assert_allclose(g, gr, rtol=0, atol=max(rel * float(np.abs(gr).max()), 1e-2 * floor, 5e-7))
            continue
        np.testing.assert_allclose(g, gr, rtol=0, atol=rel * max(float(np.abs(gr).max()), floor))


def assert_params_close(got, want, K, lr=3e-4):
    for w, wr in zip(got, want):
        d = np.abs(np.asarray(w).reshape(-1) - np.asarray(wr).reshape(-1))
        assert d.max() <= 2 * K * lr
        assert np.sum(d > 3e-5) <= max(1, 0.02 * d.size)


@pytest.mark.parametrize("name", sorted(CASES))
def test_reference_fixture(golden, name):
    c = load_case(golden, name)
    tr, net, opt = make_trainer(c["w0"], c["act"], c["pool"])
    for k in range(c["K"]):
        log = tr.train(tb(c["batches"][k]), return_grad=(k == 0))
        assert log["training_loss"] == pytest.approx(c["loss"][k], rel=1e-4)
        assert log["training_acc"] == pytest.approx(c["acc"][k], abs=1e-6)
        if k == 0:
            assert_grads_close(split(log["grad"], c["w0"]), c["g0"])
    ev = tr.evaluation(tb(c["batches"][c["K"]]))
    assert ev["eval_loss"] == pytest.approx(c["eval"][0], rel=1e-4)
    assert ev["eval_acc"] == pytest.approx(c["eval"][1], abs=1e-6)
    assert_params_close([p.detach().cpu().numpy() for p in net.parameters()], c["wK"], c["K"])
    sd = tr.state_dict()
    assert set(sd) == {"net", "optimizer"}
    st = sd["optimizer"]["state"]
    assert len(st) == len(c["w0"]) and float(st[0]["step"]) == c["K"]
    for i, (mr, vr) in enumerate(zip(c["mK"], c["vK"])):
        m = st[i]["exp_avg"].cpu().numpy(); v = st[i]["exp_avg_sq"].cpu().numpy()
        assert m.shape == mr.shape
        np.testing.assert_allclose(m, mr, rtol=2e-3, atol=2e-6 * max(1.0, float(np.abs(mr).max())))
        np.testing.assert_allclose(v, vr, rtol=4e-3, atol=1e-9)


@pytest.mark.parametrize("act", ["relu", "tanh", "cos", "sin", "linear", "rbf", "swish", "softplus", "leaky_relu"])
@pytest.mark.parametrize("pool", ["mean", "sum"])
def test_gradients_every_activation(act, pool):
    rng = np.random.default_rng(sum(map(ord, act + pool)))
    S, A, B, T = 7, 3, 33, 9
    widths = [40, 24, 16]
    dims = [S + A] + widths + [1]
    weights = []
    for i in range(len(dims) - 1):
        weights += [rng.standard_normal((dims[i], dims[i + 1])).astype(np.float32),
                    (0.1 * rng.standard_normal(dims[i + 1])).astype(np.float32)]
    batch = synth_batch(rng, B, T, S, A)
    batch[3][0, :] = 0.0  # a segment with an empty mask: n clamps to 1 (util.py:363)
    loss, acc, grads = mr_loss_and_grads(weights, batch, act, pool)
    tr, net, _ = make_trainer(weights, act, pool)
    log = tr.train(tb(batch), return_grad=True)
    assert log["training_loss"] == pytest.approx(loss, rel=1e-4)
    assert log["training_acc"] == pytest.approx(acc, abs=1e-6)
    assert_grads_close(split(log["grad"], weights), grads)


def test_nan_rewards_are_skipped_like_nansum():
    rng = np.random.default_rng(5)
    S, A, B, T = 4, 2, 8, 6
    weights = [rng.standard_normal((6, 16)).astype(np.float32), np.zeros(16, np.float32),
               rng.standard_normal((16, 1)).astype(np.float32), np.zeros(1, np.float32)]
    batch = synth_batch(rng, B, T, S, A, ragged=False)
    batch[0][2, 3, :] = np.nan  # one row of segment 1 predicts NaN
    tr, _, _ = make_trainer(weights, "relu", "mean")
    ev = tr.evaluation(tb(batch))
    assert np.isfinite(ev["eval_loss"])
    from oracle.mr_train_oracle import ce_loss_acc, mr_forward
    logits, _ = mr_forward(weights, batch, "relu", "mean")
    loss, acc, _ = ce_loss_acc(logits, batch[8])
    assert ev["eval_loss"] == pytest.approx(loss, rel=1e-4) and ev["eval_acc"] == pytest.approx(acc, abs=1e-6)


@pytest.mark.parametrize("precision", ["fp32", "split_bf16"])
def test_production_shape_against_oracle_trajectory(precision):
    """antmaze MR shape (37 -> 128 x 3 -> 1), batch 256 x 25-step segments (12 800 rows, > one chunk per SM wave).
    precision="split_bf16" (opt-in): the step's seven GEMMs on the tensor cores, three bf16 MMAs per product, i.e. 2^-17
    relative PER PRODUCT.  Loss, accuracy and the parameter bars are those of the fp32 path; the first-step gradients
    are held to 1.5e-2 of a tensor's largest entry instead of 1e-3 (measured 3e-3 .. 7e-3): a Bradley-Terry gradient near
    the initial ln 2 plateau is a sum over 12 800 rows of terms of alternating sign, sum |terms| / |sum| ~ 500, which
    amplifies the product error of the weight-gradient GEMMs (fp32 FFMA: 1e-6 on the same tensors)."""
    rng = np.random.default_rng(21)
    S, A, B, T, K = 29, 8, 256, 25, 3
    dims = [S + A, 128, 128, 128, 1]
    weights = []
    for i in range(4):
        weights += [rng.standard_normal((dims[i], dims[i + 1])).astype(np.float32), np.zeros(dims[i + 1], np.float32)]
    orc = MRTrainerOracle(weights, "relu")
    tr, net, _ = make_trainer(weights, "relu", "mean", precision=precision)
    for k in range(K):
        batch = synth_batch(rng, B, T, S, A)
        want = orc.train(batch)
        got = tr.train(tb(batch), return_grad=(k == 0))
        if k == 0:  # first-step gradients, tensor by tensor, at 1e-3 of the tensor's largest entry
            _, _, g_ref = mr_loss_and_grads(weights, batch, "relu", "mean")
            flat = got["grad"].cpu().numpy()
            off = 0
            for g in g_ref:
                part = flat[off:off + g.size].reshape(g.shape)
                off += g.size
                if g.size > 1:  # (the output bias gradient is a cancelling sum: exactly 0 in real arithmetic)
                    rel = 1e-3 if precision == "fp32" else 1.5e-2
                    assert np.abs(part - g).max() <= rel * np.abs(g).max() + 1e-9
        assert got["training_loss"] == pytest.approx(want["training_loss"], rel=2e-4)
        assert abs(got["training_acc"] - want["training_acc"]) <= 1.5 / B  # a pair with |logit0 - logit1| ~ 1e-7
    assert_params_close([p.detach().cpu().numpy() for p in net.parameters()], orc.weights, K)


def test_trained_net_feeds_the_relabel_path_and_checkpoints_round_trip(tmp_path):
    from iqlpref_b200.relabel import ensemble_predict
    rng = np.random.default_rng(3)
    S, A, B, T = 5, 2, 12, 4
    weights = [rng.standard_normal((7, 64)).astype(np.float32), np.zeros(64, np.float32),
               rng.standard_normal((64, 1)).astype(np.float32), np.zeros(1, np.float32)]
    tr, net, opt = make_trainer(weights, "tanh", "mean")
    b0, b1 = synth_batch(rng, B, T, S, A), synth_batch(rng, B, T, S, A)
    tr.train(tb(b0))
    torch.save(tr.state_dict(), tmp_path / "checkpoint_0.pt")  # run_mr_training.py:147-151
    after1 = tr.train(tb(b1))
    x = torch.from_numpy(rng.standard_normal((50, 7)).astype(np.float32)).cuda()
    r_direct = ensemble_predict([tr.member()], x, "mean", activation="tanh")
    assert torch.equal(r_direct, net(x).reshape(-1))
    # resume from the file in a fresh trainer: identical continuation
    # sync=False: same step, the log holds device scalars
    tr3, net3, _ = make_trainer(weights, "tanh", "mean")
    tr3.train(tb(b0), sync=False)
    lazy = tr3.train([t.cuda() for t in tb(b1)], sync=False)
    assert torch.is_tensor(lazy["training_loss"]) and float(lazy["training_loss"]) == after1["training_loss"]
    tr2, net2, _ = make_trainer(weights, "tanh", "mean")
    tr2.load_state_dict(torch.load(tmp_path / "checkpoint_0.pt", weights_only=False))
    assert tr2.step_count == 1
    again = tr2.train(tb(b1))
    assert again == after1
    for p, q in zip(net.parameters(), net2.parameters()):
        assert torch.equal(p, q)


def test_errors():
    from iqlpref_b200.mr_training import MRTrainer
    from iqlpref_b200.reward_models.q_mlp import MLP
    net = MLP(6, 1, [8], "relu")
    opt = torch.optim.Adam(net.parameters(), lr=1e-3)
    with pytest.raises(ValueError):
        MRTrainer(net, opt, 1, device="cuda", bt_pool="median")
    with pytest.raises(RuntimeError):
        MRTrainer(net, opt, 1, device="cpu")
    tr = MRTrainer(net, opt, 1, device="cuda")
    bad = synth_batch(np.random.default_rng(0), 2, 3, 5, 2)  # 7 inputs for a 6-input net
    with pytest.raises(RuntimeError):
        tr.train(tb(bad))


def test_fit_loop_emits_the_snapshot_ensemble_the_relabel_reads(tmp_path):
    """run_mr_training.py:117-215 -> checkpoint_{epoch}.pt / best_model.pt -> qlearning_dataset_mr_ensemble (iql.py:1085)."""
    from iqlpref_b200.mr_training import fit_mr
    from iqlpref_b200.relabel import _discover_mr_snapshots, ensemble_predict, qlearning_dataset_mr_ensemble, _state_to_member
    rng = np.random.default_rng(8)
    S, A, B, T = 6, 2, 16, 5
    weights = [rng.standard_normal((8, 64)).astype(np.float32), np.zeros(64, np.float32),
               rng.standard_normal((64, 64)).astype(np.float32), np.zeros(64, np.float32),
               rng.standard_normal((64, 1)).astype(np.float32), np.zeros(1, np.float32)]
    tr, net, _ = make_trainer(weights, "tanh", "mean", lr=1e-2)
    train_b = [tb(synth_batch(rng, B, T, S, A)) for _ in range(3)]
    val_b = [tb(synth_batch(rng, B, T, S, A)) for _ in range(2)]
    ckpt_dir = str(tmp_path / "mr")
    hist = fit_mr(tr, train_b, val_b, epochs=6, eval_every=2, checkpoints_path=ckpt_dir, criteria_key="loss",
                  activations="tanh")
    assert len(hist) == 7 and np.isnan(hist[0]["training_loss"]) and "eval_loss" in hist[0] and "eval_loss" not in hist[1]
    assert hist[3]["training_loss"] < hist[1]["training_loss"]  # it learns the training pairs
    paths = _discover_mr_snapshots(ckpt_dir, burn_in=2)
    assert [os.path.basename(p) for p in paths] == ["checkpoint_2.pt", "checkpoint_4.pt", "checkpoint_6.pt"]
    assert os.path.isfile(os.path.join(ckpt_dir, "best_model.pt"))
    best = torch.load(os.path.join(ckpt_dir, "best_model.pt"), weights_only=False)
    assert set(best) == {"net", "optimizer"} and "layers.linear_1.W" in best["net"]
    # the emitted directory is a valid input of the ensemble relabel; its rewards are the CVaR over the snapshots
    n = 40
    ds = {"observations": rng.standard_normal((n, S)).astype(np.float32),
          "actions": rng.uniform(-1, 1, (n, A)).astype(np.float32),
          "rewards": np.zeros(n, np.float32), "terminals": np.zeros(n, bool), "timeouts": np.zeros(n, bool)}

    class Env:
        _max_episode_steps = 1000
    with pytest.warns(RuntimeWarning):  # three snapshots: thin tail
        out = qlearning_dataset_mr_ensemble(Env(), ckpt_dir, alpha=0.5, burn_in=2, device="cuda", dataset=ds)
    members = [_state_to_member(torch.load(p, weights_only=False)["net"]) for p in paths]
    x = torch.from_numpy(np.concatenate([ds["observations"][:-1], ds["actions"][:-1]], 1)).cuda()
    want = ensemble_predict(members, x, "cvar", 0.5, activation="tanh").cpu().numpy()
    np.testing.assert_allclose(out["rewards"].reshape(-1), want, rtol=1e-6, atol=1e-7)
    # the last snapshot is the trainer's current net
    for w, p in zip(members[-1], net.parameters()):
        assert torch.equal(w.cuda(), p.detach())


@pytest.mark.parametrize("S,A,B,T,widths", [
    (3, 1, 1, 1, [5]),              # one pair, one step, one narrow layer (32-wide GEMM tiles, scalar epilogue)
    (7, 2, 5, 37, [30, 17]),        # segments longer than a warp, widths not a multiple of 4
    (11, 4, 300, 3, [132, 64, 8]),  # more pairs than one reduction block, a width just over one tile
    (29, 8, 67, 41, [128, 128, 128, 128]),  # 5 494 rows: row chunks with a ragged tail, four hidden layers
    (2, 0, 9, 4, [16]),             # action_dim 0
])
def test_odd_shapes_against_oracle(S, A, B, T, widths):
    rng = np.random.default_rng(S * 1000 + B)
    dims = [S + A] + widths + [1]
    weights = []
    for i in range(len(dims) - 1):
        weights += [rng.standard_normal((dims[i], dims[i + 1])).astype(np.float32),
                    (0.1 * rng.standard_normal(dims[i + 1])).astype(np.float32)]
    batch = synth_batch(rng, B, T, S, A)
    loss, acc, grads = mr_loss_and_grads(weights, batch, "swish", "mean")
    tr, net, _ = make_trainer(weights, "swish", "mean")
    log = tr.train(tb(batch), return_grad=True)
    assert log["training_loss"] == pytest.approx(loss, rel=1e-4)
    assert abs(log["training_acc"] - acc) <= 1.01 / B
    assert_grads_close(split(log["grad"], weights), grads)
    # a second, smaller batch through the same handle (row count changes: the task table is rebuilt)
    if B > 1:
        small = tuple(x[: B // 2 + 1] for x in batch)
        loss2, _, grads2 = mr_loss_and_grads([p.detach().cpu().numpy() for p in net.parameters()], small, "swish", "mean")
        log2 = tr.train(tb(small), return_grad=True)
        assert log2["training_loss"] == pytest.approx(loss2, rel=1e-4)
        assert_grads_close(split(log2["grad"], weights), grads2)
