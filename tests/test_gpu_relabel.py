"""GPU parity: reward-model relabelling (MLP, ensembles with mean/CVaR, BNN tuples) against fixtures
produced by the unmodified reference.  Predictions: 1e-3 relative (observed ~1e-6); the CVaR member
selection must be bit-exact given the predictions."""
import os
import tempfile
import warnings

import numpy as np
import pytest
import torch
import yaml

from oracle import relabel_oracle as ro

pytestmark = pytest.mark.gpu


class FakeEnv:
    _max_episode_steps = 60


@pytest.fixture(scope="module")
def rel(golden_dir):
    return np.load(os.path.join(golden_dir, "relabel.npz"))


def _ds(rel, tag):
    d = {k: rel[f"{tag}.{k}"] for k in ("observations", "actions", "rewards", "terminals")}
    if f"{tag}.timeouts" in rel.files:
        d["timeouts"] = rel[f"{tag}.timeouts"]
    return d


@pytest.mark.parametrize("act", ["relu", "tanh", "swish", "softplus", "leaky_relu", "rbf", "sin", "cos", "linear"])
def test_reward_mlp_forward_all_activations(rel, act):
    from iqlpref_b200.relabel import ensemble_predict
    w = [rel[f"mlp.{act}.w{i}"] for i in range(8)]
    y = ensemble_predict([w], rel["mlp.x"], "mean", activation=act).cpu().numpy()
    np.testing.assert_allclose(y, rel[f"mlp.{act}.y"][:, 0], rtol=1e-3, atol=1e-5)


def test_reward_mlp_module_surface(rel):
    from iqlpref_b200.reward_models import MLP, Q_MLP
    m = MLP(37, 1, [32, 48, 32], "tanh").cuda()
    m.load_state_dict({k: torch.from_numpy(rel[f"mlp.tanh.w{i}"]) for i, k in enumerate(m.state_dict())})
    x = torch.from_numpy(rel["mlp.x"]).cuda()
    y = m(x)
    assert tuple(y.shape) == (200, 1)
    np.testing.assert_allclose(y.cpu().numpy(), rel["mlp.tanh.y"], rtol=1e-3, atol=1e-5)
    np.testing.assert_allclose(m.predict(x).cpu().numpy(), rel["mlp.tanh.y"], rtol=1e-3, atol=1e-5)
    q = Q_MLP(37, 1, [32, 48, 32], "tanh").cuda()
    q.load_state_dict(m.state_dict())
    np.testing.assert_allclose(q(x[:, :29], x[:, 29:]).cpu().numpy(), rel["mlp.tanh.y"], rtol=1e-3, atol=1e-5)


@pytest.mark.parametrize("tag", ["ds_t", "ds_n"])
@pytest.mark.parametrize("toe", [False, True])
def test_qlearning_dataset_mr(rel, tag, toe):
    from iqlpref_b200.iql import qlearning_dataset_mr
    from iqlpref_b200.reward_models import MLP
    m = MLP(37, 1, [32, 32, 32], "relu").cuda()
    m.load_state_dict({k: torch.from_numpy(rel[f"mr.w{i}"]) for i, k in enumerate(m.state_dict())})
    out = qlearning_dataset_mr(FakeEnv(), m, dataset=_ds(rel, tag), terminate_on_end=toe)
    pre = f"mr.{tag}.toe{int(toe)}."
    for k in ("observations", "actions", "next_observations", "terminals"):
        np.testing.assert_array_equal(out[k], rel[pre + k])
    np.testing.assert_allclose(out["rewards"], rel[pre + "rewards"], rtol=1e-3, atol=1e-5)


@pytest.mark.parametrize("alpha", [0.0, 0.5, 0.9, 0.95])
@pytest.mark.parametrize("burn", [0, 10])
def test_qlearning_dataset_mr_ensemble(rel, alpha, burn):
    from iqlpref_b200.iql import qlearning_dataset_mr_ensemble
    from iqlpref_b200.reward_models import MLP
    with tempfile.TemporaryDirectory() as td:
        with open(os.path.join(td, "config.yaml"), "w") as f:
            yaml.safe_dump({"activations": "tanh"}, f)
        for e in range(0, 60, 5):
            m = MLP(37, 1, [32, 32], "tanh")
            sd = {k: torch.from_numpy(rel[f"ens.e{e}.w{i}"]) for i, k in enumerate(m.state_dict())}
            if e % 10 == 0:
                sd = {"_orig_mod." + k: v for k, v in sd.items()}
            torch.save({"net": sd}, os.path.join(td, f"checkpoint_{e}.pt"))
        torch.save({"net": sd}, os.path.join(td, "best_model.pt"))  # must be ignored (iql.py:1050-1052)
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            out = qlearning_dataset_mr_ensemble(FakeEnv(), td, alpha=alpha, burn_in=burn, device="cuda",
                                                dataset=_ds(rel, "ds_t"))
        np.testing.assert_allclose(out["rewards"], rel[f"ens.a{alpha}.b{burn}.rewards"], rtol=1e-3, atol=1e-5)
        with pytest.raises(ValueError):
            qlearning_dataset_mr_ensemble(FakeEnv(), td, alpha=1.0, dataset=_ds(rel, "ds_t"))
        with pytest.raises(ValueError):
            qlearning_dataset_mr_ensemble(FakeEnv(), td, alpha=0.5, burn_in=1000, dataset=_ds(rel, "ds_t"))
    with tempfile.TemporaryDirectory() as td, pytest.raises(FileNotFoundError):
        open(os.path.join(td, "config.yaml"), "w").write("activations: relu\n")
        qlearning_dataset_mr_ensemble(FakeEnv(), td, alpha=0.5, dataset=_ds(rel, "ds_t"))


@pytest.mark.parametrize("alpha,ns", [(0.9, 500), (0.9, 16), (0.0, 40), (0.75, 40)])
def test_qlearning_dataset_bnn(rel, alpha, ns):
    from iqlpref_b200.iql import qlearning_dataset_bnn
    with tempfile.TemporaryDirectory() as td:
        for c in range(2):
            d = os.path.join(td, "sampling_f", f"chain_{c}", "sampled_weights")
            os.makedirs(d)
            samples = [tuple(rel[f"bnn.c{c}.k{k}.w{i}"] for i in range(8)) for k in range(20)]
            torch.save({"sampled_weights": samples}, os.path.join(d, "sampled_weights_0000000"))
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            out = qlearning_dataset_bnn(FakeEnv(), td, alpha=alpha, n_samples=ns, device="cuda", dataset=_ds(rel, "ds_t"))
        np.testing.assert_allclose(out["rewards"], rel[f"bnn.a{alpha}.n{ns}.rewards"], rtol=1e-3, atol=1e-5)
    with tempfile.TemporaryDirectory() as td, pytest.raises(FileNotFoundError):
        qlearning_dataset_bnn(FakeEnv(), td, alpha=0.5, dataset=_ds(rel, "ds_t"))


@pytest.mark.parametrize("S,alpha", [(40, 0.9), (40, 0.5), (10, 0.9), (981, 0.95), (7, 0.0), (1, 0.5)])
def test_cvar_selection_mask_is_bit_exact(S, alpha):
    """The tail membership must equal the oracle's on the SAME predictions, ties included."""
    from iqlpref_b200.relabel import ensemble_predict
    rng = np.random.default_rng(S)
    d_in, n = 13, 3000
    members = [[rng.standard_normal((d_in, 16)).astype(np.float32), rng.standard_normal(16).astype(np.float32),
                rng.standard_normal((16, 1)).astype(np.float32), rng.standard_normal(1).astype(np.float32)]
               for _ in range(S)]
    if S >= 7:  # exact duplicates -> ties at the threshold
        members[3] = members[1]
        members[5] = members[1]
    x = rng.standard_normal((n, d_in)).astype(np.float32)
    out, preds, sel = ensemble_predict(members, x, "cvar", alpha, "relu", return_preds=True, return_selection=True)
    preds_h, sel_h = preds.cpu().numpy(), sel.cpu().numpy().astype(bool)
    np.testing.assert_array_equal(sel_h, ro.cvar_selection(preds_h, alpha))
    np.testing.assert_allclose(out.cpu().numpy(), ro.cvar_reduce(preds_h, alpha), rtol=1e-5, atol=1e-6)
    # and the predictions themselves agree with the oracle's forward
    want = np.stack([ro.reward_mlp_forward(x, m, "relu")[:, 0] for m in members[:5]])
    np.testing.assert_allclose(preds_h[:5], want, rtol=1e-3, atol=1e-5)


def test_relabel_large_ragged_rows():
    """Row counts that are not tile multiples, > one chunk, and the empty input."""
    from iqlpref_b200.relabel import ensemble_predict
    rng = np.random.default_rng(3)
    members = [[rng.standard_normal((37, 128)).astype(np.float32), np.zeros(128, np.float32),
                rng.standard_normal((128, 128)).astype(np.float32), np.zeros(128, np.float32),
                rng.standard_normal((128, 1)).astype(np.float32), np.zeros(1, np.float32)] for _ in range(3)]
    for n in (1, 127, 129, 70_001):
        x = rng.standard_normal((n, 37)).astype(np.float32)
        got = ensemble_predict(members, x, "mean", activation="relu").cpu().numpy()
        want = np.mean([ro.reward_mlp_forward(x, m, "relu")[:, 0] for m in members], axis=0)
        np.testing.assert_allclose(got, want, rtol=1e-3, atol=1e-4)
    assert ensemble_predict(members, np.zeros((0, 37), np.float32), "mean").shape == (0,)
    with pytest.raises(ValueError):
        ensemble_predict(members, np.zeros((4, 36), np.float32), "mean")
    with pytest.raises(ValueError):
        ensemble_predict(members, np.zeros((4, 37), np.float32), "cvar", alpha=1.0)


def _members(rng, d_in, w, depth, S, bias=True):
    out = []
    for _ in range(S):
        m, n_in = [], d_in
        for _l in range(depth):
            m += [rng.standard_normal((n_in, w)).astype(np.float32),
                  (rng.standard_normal(w) if bias else np.zeros(w)).astype(np.float32)]
            n_in = w
        m += [rng.standard_normal((w, 1)).astype(np.float32), rng.standard_normal(1).astype(np.float32)]
        out.append(m)
    return out


# split-bf16 tolerance: each product carries ~2^-18..2^-16 relative error (three bf16 MMAs, fp32 accumulation), so a
# prediction is within 1e-3 relative of the reference, with an absolute floor of 1e-4 of the prediction scale.
def _assert_split_bf16_close(got, want):
    np.testing.assert_allclose(got, want, rtol=1e-3, atol=1e-4 * max(1.0, float(np.abs(want).max())))


@pytest.mark.parametrize("d_in,w,depth,S,act,n", [
    (37, 128, 3, 1, "relu", 1000),      # antmaze q_mlp (BASELINE configs[1])
    (37, 128, 3, 1, "relu", 70_001),    # ragged, more tiles than SMs
    (37, 128, 3, 7, "relu", 40_000),    # ensemble, every CTA streams all members (weight slots reloaded)
    (69, 64, 3, 200, "relu", 300),      # pen-human BNN: members split across CTAs, 2-3 members per CTA
    (69, 128, 3, 10, "tanh", 5000),     # pen-human snapshot ensemble, generic activation
    (128, 128, 2, 3, "swish", 513),
    (16, 64, 1, 3, "relu", 257),
    (5, 64, 2, 1, "leaky_relu", 1),
])
def test_tensor_core_relabel_matches_oracle_and_fp32_path(d_in, w, depth, S, act, n):
    from iqlpref_b200.relabel import ensemble_predict, pack_members
    rng = np.random.default_rng(d_in * 1000 + w + S)
    members = _members(rng, d_in, w, depth, S)
    x = rng.standard_normal((n, d_in)).astype(np.float32)
    packed = pack_members(members)
    out_t, preds_t = ensemble_predict(packed, x, "cvar", 0.5, act, return_preds=True, precision="split_bf16")
    out_f, preds_f = ensemble_predict(packed, x, "cvar", 0.5, act, return_preds=True, precision="fp32")
    preds_t, preds_f = preds_t.cpu().numpy(), preds_f.cpu().numpy()
    _assert_split_bf16_close(preds_t, preds_f)
    for s in list(range(min(S, 3))) + [S - 1]:
        _assert_split_bf16_close(preds_t[s], ro.reward_mlp_forward(x, members[s], act)[:, 0])
    # the reduction is exact given the predictions, whatever produced them
    np.testing.assert_allclose(out_t.cpu().numpy(), ro.cvar_reduce(preds_t, 0.5), rtol=1e-5, atol=1e-6)
    # without the side outputs (scratch predictions, or the single-model direct write) the result is the same
    out_only = ensemble_predict(packed, x, "cvar", 0.5, act, precision="split_bf16")
    np.testing.assert_array_equal(out_only.cpu().numpy(), out_t.cpu().numpy())


def test_tensor_core_relabel_reads_row_strided_tables():
    """The relabel reads [obs|act] straight out of the packed replay table (row stride 68 floats)."""
    from iqlpref_b200.relabel import ensemble_predict
    rng = np.random.default_rng(11)
    members = _members(rng, 37, 128, 3, 1)
    table = torch.from_numpy(rng.standard_normal((3000, 68)).astype(np.float32)).cuda()
    for view in (table[:, :37], table[:, 1:38]):  # 16-byte aligned rows and not
        got = ensemble_predict(members, view, "mean", precision="split_bf16").cpu().numpy()
        _assert_split_bf16_close(got, ro.reward_mlp_forward(view.cpu().numpy(), members[0], "relu")[:, 0])


def test_tensor_core_relabel_refuses_other_shapes():
    from iqlpref_b200.relabel import ensemble_predict
    rng = np.random.default_rng(12)
    x = rng.standard_normal((64, 37)).astype(np.float32)
    with pytest.raises(RuntimeError):
        ensemble_predict(_members(rng, 37, 96, 2, 1), x, "mean", precision="split_bf16")
    got = ensemble_predict(_members(np.random.default_rng(12), 37, 96, 2, 1), x[:, :37], "mean", precision="auto")
    assert got.shape == (64,)
    with pytest.raises(ValueError):
        ensemble_predict(_members(rng, 37, 128, 2, 1), x, "mean", precision="tf32")
