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


# ------------------------------------------------------------------------------------------------
# CVaR tail membership against the REFERENCE's fp32 predictions, default precision included
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("precision,rel_tol", [("auto", 4e-5), ("fp32", 4e-6)])
@pytest.mark.parametrize("alpha", [0.5, 0.9, 0.95])
@pytest.mark.parametrize("S", [40, 981])
def test_cvar_tail_membership_vs_reference_fp32_predictions(golden_dir, S, alpha, precision, rel_tol):
    """`tests/golden/cvar_ref.npz` holds the all_preds matrix the unmodified reference RewardMLP computes in fp32 on the
    CPU (oracle/gen_golden_cvar.py).  The reference's tail set of a row is `np.argpartition(all_preds, kth, axis=0)
    [:n_tail]` (iql.py:1009-1011, 1185-1187).  The CUDA path selects from its own predictions (default precision "auto" =
    split-bf16 on the tensor cores, ~1e-5 of the prediction scale; "fp32" = FFMA), so membership must be IDENTICAL
    wherever the reference separates the n_tail-th and (n_tail+1)-th smallest member by more than twice the arithmetic
    tolerance, and may differ only for members that close to the threshold (the tie-tolerant criterion of SURVEY H6)."""
    from iqlpref_b200.relabel import cvar_n_tail, ensemble_predict
    from oracle.gen_golden_cvar import members_from_seed, rows_from_seed
    ref = np.load(os.path.join(golden_dir, "cvar_ref.npz"))[f"preds.S{S}"]
    x = rows_from_seed()
    out, preds, sel = ensemble_predict(members_from_seed(S, seed=S), x, "cvar", alpha, "relu", return_preds=True,
                                       return_selection=True, precision=precision)
    preds, sel, out = preds.cpu().numpy(), sel.cpu().numpy().astype(bool), out.cpu().numpy()
    n_tail = cvar_n_tail(alpha, S)
    assert n_tail == max(1, int(np.floor((1.0 - alpha) * S)))
    kth = min(n_tail, S - 1)
    ref_sel = np.zeros_like(sel)
    np.put_along_axis(ref_sel, np.argpartition(ref, kth, axis=0)[:n_tail], True, axis=0)
    tol = rel_tol * float(np.abs(ref).max())
    err = float(np.abs(preds - ref).max())
    srt = np.sort(ref, axis=0)
    clear = (srt[n_tail] - srt[n_tail - 1]) > 2 * tol
    diff = sel != ref_sel
    print(f"S={S} alpha={alpha} {precision}: max |pred - ref| = {err:.2e} (tol {tol:.2e}); rows with a clear threshold "
          f"{int(clear.sum())}/{clear.size}; rows whose tail set differs {int(diff.any(axis=0).sum())}")
    assert err <= tol
    assert (sel.sum(axis=0) == n_tail).all()
    assert clear.mean() > 0.4
    np.testing.assert_array_equal(sel[:, clear], ref_sel[:, clear])
    # where membership differs, the member sits within the tolerance band around the reference's threshold
    thr = 0.5 * (srt[n_tail] + srt[n_tail - 1])
    assert (np.abs(ref - thr[None, :])[diff] <= 2 * tol).all()
    want = np.partition(ref, kth, axis=0)[:n_tail].mean(axis=0)
    np.testing.assert_allclose(out, want, rtol=0, atol=2 * tol)


def test_empirical_cvar_matches_the_reference(rel):
    """product `empirical_cvar` (iql.py:735-763) against values written by the reference function (relabel.npz)."""
    from iqlpref_b200.iql import empirical_cvar
    got = [empirical_cvar(rel["cvar.samples"], a) for a in (0.0, 0.5, 0.9, 0.95, 0.99)]
    np.testing.assert_allclose(got, rel["cvar.values"], rtol=1e-6)
    assert abs(got[0] - float(rel["cvar.samples"].mean())) < 1e-6  # alpha = 0 is the plain mean (docstring :751-752)
    for bad in (-0.1, 1.0, 1.5):
        with pytest.raises(ValueError):
            empirical_cvar(rel["cvar.samples"], bad)


@pytest.mark.parametrize("compiled", [False, True])
def test_load_mlp_reward_model(rel, tmp_path, compiled):
    """`load_mlp_reward_model` (iql.py:1345-1353): best_model.pt["net"] + config.yaml "activations"; dims inferred from
    the shapes (:1326-1336); `_orig_mod.` prefixes of a torch.compile'd model stripped (:1312-1323)."""
    from iqlpref_b200.iql import load_mlp_reward_model, qlearning_dataset_mr
    from iqlpref_b200.reward_models import MLP
    m = MLP(37, 1, [32, 48, 32], "tanh")
    sd = {k: torch.from_numpy(rel[f"mlp.tanh.w{i}"]) for i, k in enumerate(m.state_dict())}
    if compiled:
        sd = {"_orig_mod." + k: v for k, v in sd.items()}
    torch.save({"net": sd, "optimizer": {}}, tmp_path / "best_model.pt")
    (tmp_path / "config.yaml").write_text(yaml.safe_dump({"activations": "tanh", "hidden_dims": [1, 2, 3]}))
    model = load_mlp_reward_model(str(tmp_path), "cuda")
    assert not model.training and model.hidden_dims == [32, 48, 32] and model.input_dim == 37
    assert model.activation_name == "tanh" and next(model.parameters()).is_cuda
    y = model(torch.from_numpy(rel["mlp.x"]).cuda())
    np.testing.assert_allclose(y.cpu().numpy(), rel["mlp.tanh.y"], rtol=1e-3, atol=1e-5)
    # config.yaml without "activations" means relu (iql.py:1339-1342)
    m2 = MLP(37, 1, [32, 32, 32], "relu")
    torch.save({"net": {k: torch.from_numpy(rel[f"mr.w{i}"]) for i, k in enumerate(m2.state_dict())}},
               tmp_path / "best_model.pt")
    (tmp_path / "config.yaml").write_text("lr: 0.001\n")
    model = load_mlp_reward_model(str(tmp_path), "cuda")
    assert model.activation_name == "relu"
    out = qlearning_dataset_mr(FakeEnv(), model, dataset=_ds(rel, "ds_t"))
    np.testing.assert_allclose(out["rewards"], rel["mr.ds_t.toe0.rewards"], rtol=1e-3, atol=1e-5)
    with pytest.raises(FileNotFoundError):
        load_mlp_reward_model(str(tmp_path / "nowhere"), "cuda")


@pytest.mark.parametrize("L", [1, 2])
def test_load_pt_reward_model(rel, tmp_path, L):
    """`load_pt_reward_model` (iql.py:1356-1390): every dimension inferred from the state-dict shapes, heads / eps /
    intermediate_dim from config.yaml; the loaded model relabels like the reference's."""
    from iqlpref_b200.iql import load_pt_reward_model, qlearning_dataset_pt
    pre = f"pt{L}."
    sd = {k[len(pre):]: torch.from_numpy(rel[k]) for k in rel.files
          if k.startswith(pre) and ".fw." not in k and ".ds_" not in k}
    max_pos = 64
    for l in range(L):  # the reference's state_dict carries the causal buffer (its shape gives max_pos, :1386)
        sd[f"gpt.layers.{l}.attention.causal_bias"] = torch.triu(torch.full((max_pos, max_pos), float("-inf")),
                                                                 diagonal=1).view(1, 1, max_pos, max_pos)
    torch.save({"net": {"_orig_mod." + k: v for k, v in sd.items()}}, tmp_path / "best_model.pt")
    (tmp_path / "config.yaml").write_text(yaml.safe_dump({"num_heads": 4, "intermediate_dim": 64, "model_eps": 0.1}))
    model = load_pt_reward_model(str(tmp_path), "cuda")
    assert not model.training and model.num_layers == L and model.embd_dim == 32 and model.pref_attn_embd_dim == 32
    assert model.timestep_embed.weight.shape[0] == 61 and abs(model.stacked_layer_norm.eps - 0.1) < 1e-12
    out = qlearning_dataset_pt(FakeEnv(), model, query_length=10, dataset=_ds(rel, "ds_t"))
    np.testing.assert_allclose(out["rewards"], rel[f"pt{L}.ds_t.rewards"], rtol=1e-3, atol=1e-5)


def test_ensemble_diagnostics_without_the_prediction_matrix(capsys):
    """`_ensemble_rewards` (the CVaR reward + the reference's "[CVaR stability]" and "ensemble mean +- std" prints,
    iql.py:1013-1021, 1188-1204) evaluates the stability check on its <= 50 transitions only and takes the member mean
    out of the reduction pass: the numbers must be the ones the full [S, N] matrix gives."""
    from iqlpref_b200 import relabel
    rng = np.random.default_rng(17)
    S, n, d_in = 40, 5000, 37
    members = _members(rng, d_in, 64, 3, S)
    x = rng.standard_normal((n, d_in)).astype(np.float32)
    out_full, preds = relabel.ensemble_predict(members, x, "cvar", 0.9, "relu", return_preds=True)
    out_lean, mean = relabel.ensemble_predict(members, x, "cvar", 0.9, "relu", return_mean=True)
    assert torch.equal(out_full, out_lean)
    np.testing.assert_allclose(mean.cpu().numpy(), preds.double().mean(0).cpu().numpy(), rtol=1e-6, atol=1e-7)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        capsys.readouterr()
        want_ratio = relabel.cvar_stability_check(preds, 0.9)
        capsys.readouterr()
        got = relabel._ensemble_rewards(members, x, 0.9, "relu", "cuda", "MR/CVaR", "remedy", "Ensemble mean reward")
        printed = capsys.readouterr().out
    np.testing.assert_array_equal(got, out_full.cpu().numpy())
    pm = preds.mean(0)
    assert f"mean relative diff = {want_ratio:.3f}" in printed
    assert f"Ensemble mean reward: {pm.mean().item():.4f} ± {pm.std(unbiased=False).item():.4f}" in printed
    # alpha = 0: the check is skipped like the reference's, the reward is the plain mean
    got0 = relabel._ensemble_rewards(members, x, 0.0, "relu", "cuda", "MR/CVaR", "remedy", "Ensemble mean reward")
    assert "stability check skipped" in capsys.readouterr().out
    np.testing.assert_allclose(got0, pm.cpu().numpy(), rtol=1e-5, atol=1e-6)


def test_large_ensemble_streams_in_row_chunks():
    """S = 981 members x 400 000 rows: the [S, N] matrix would be 1.6 GB; the library walks row chunks of a <= 512 MB
    scratch and the result equals a call that covers a slice of the rows on its own."""
    from iqlpref_b200.relabel import ensemble_predict, pack_members
    rng = np.random.default_rng(23)
    packed = pack_members(_members(rng, 69, 64, 3, 981))
    x = torch.randn((400_000, 69), device="cuda", generator=torch.Generator(device="cuda").manual_seed(1))
    free0, _ = torch.cuda.mem_get_info()
    out, mean = ensemble_predict(packed, x, "cvar", 0.95, "relu", return_mean=True)
    torch.cuda.synchronize()
    part = ensemble_predict(packed, x[250_000:251_000], "cvar", 0.95, "relu")
    assert torch.equal(out[250_000:251_000], part)
    assert torch.isfinite(out).all() and torch.isfinite(mean).all() and (out <= mean + 1e-6).all()
