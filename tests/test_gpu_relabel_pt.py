"""GPU parity: preference-transformer forward and relabel (iqlb_pt_forward / iqlb_relabel_pt) against
fixtures produced by the unmodified reference (RewardPT.forward, qlearning_dataset_pt) and against the
numpy oracle on larger shapes.  fp32 arithmetic: tolerance 1e-3 relative (observed ~1e-5)."""
import os

import numpy as np
import pytest
import torch

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


def _model(rel, L):
    from iqlpref_b200.iql import RewardPT
    pre = f"pt{L}."
    sd = {k[len(pre):]: torch.from_numpy(rel[k]) for k in rel.files if k.startswith(pre) and ".fw." not in k and ".ds_" not in k}
    S, A = sd["state_linear.weight"].shape[1], sd["action_linear.weight"].shape[1]
    m = RewardPT(state_dim=S, action_dim=A, max_episode_steps=60, embd_dim=32, pref_attn_embd_dim=32, num_heads=4,
                 intermediate_dim=64, num_layers=L, max_pos=64, eps=0.1)
    m.load_state_dict(sd, strict=False)
    return m.cuda().eval()


@pytest.mark.parametrize("L", [1, 2])
def test_pt_forward_matches_reference(rel, L):
    m = _model(rel, L)
    res, attn = m(torch.from_numpy(rel[f"pt{L}.fw.sts"]).cuda(), torch.from_numpy(rel[f"pt{L}.fw.acts"]).cuda(),
                  torch.from_numpy(rel[f"pt{L}.fw.ts"]).cuda(), torch.from_numpy(rel[f"pt{L}.fw.am"]).cuda())
    got = res["value"][:, 0, :, 0].cpu().numpy()
    assert res["value"].shape == (6, 1, 10, 1) and len(attn) == L + 1 and all(a is None for a in attn)
    np.testing.assert_allclose(got, rel[f"pt{L}.fw.value"], rtol=1e-3, atol=1e-5)
    # the preference-attention head (pref_trans.py:303-325): what PTTrainer's loss pools (training.py:348)
    assert res["weighted_sum"].shape == (6, 10, 1)
    np.testing.assert_allclose(res["weighted_sum"][:, :, 0].cpu().numpy(), rel[f"pt{L}.fw.weighted_sum"], rtol=1e-3, atol=1e-5)


def test_pt_weighted_sum_larger_shape_vs_oracle():
    """E = P = 256, T = 100 (the antmaze PT config), ragged left padding: value and weighted_sum of every position."""
    from iqlpref_b200.iql import RewardPT
    S, A, T, n = 29, 8, 100, 9
    torch.manual_seed(4)
    m = RewardPT(state_dim=S, action_dim=A, max_episode_steps=150, embd_dim=256, pref_attn_embd_dim=256, num_heads=4,
                 intermediate_dim=1024, num_layers=1, max_pos=256, eps=0.1)
    with torch.no_grad():
        for n_, p_ in m.named_parameters():
            if "layer_norm" in n_ or n_.endswith("bias"):
                p_.add_(torch.randn_like(p_) * 0.2)
    rng = np.random.default_rng(9)
    sts = rng.standard_normal((n, T, S)).astype(np.float32)
    acts = rng.uniform(-1, 1, (n, T, A)).astype(np.float32)
    ts = np.tile(np.arange(T), (n, 1)).astype(np.int64)
    am = np.ones((n, T), np.float32)
    for j, padn in enumerate((0, 3, 99, 50, 0, 7, 1, 64, 20)):
        sts[j, :padn] = 0; acts[j, :padn] = 0; am[j, :padn] = 0
        ts[j, :padn] = 0; ts[j, padn:] = np.arange(T - padn)
    res, _ = m.cuda().eval()(torch.from_numpy(sts).cuda(), torch.from_numpy(acts).cuda(), torch.from_numpy(ts).cuda(),
                             torch.from_numpy(am).cuda())
    p = {k: v.detach().cpu().numpy() for k, v in m.state_dict().items()}
    want_v, want_w = ro.pt_forward_value(p, sts, acts, ts, am, 4, 0.1, with_weighted_sum=True)
    np.testing.assert_allclose(res["value"][:, 0, :, 0].cpu().numpy(), want_v, rtol=1e-3, atol=2e-5)
    np.testing.assert_allclose(res["weighted_sum"][:, :, 0].cpu().numpy(), want_w, rtol=1e-3, atol=2e-5)


@pytest.mark.parametrize("L", [1, 2])
@pytest.mark.parametrize("tag", ["ds_t", "ds_n"])
def test_qlearning_dataset_pt_matches_reference(rel, L, tag):
    from iqlpref_b200.iql import qlearning_dataset_pt
    ds = _ds(rel, tag)
    out = qlearning_dataset_pt(FakeEnv(), _model(rel, L), query_length=10, dataset=dict(ds))
    np.testing.assert_allclose(out["rewards"], rel[f"pt{L}.{tag}.rewards"], rtol=1e-3, atol=1e-5)
    keep, _ = ro.keep_mask_scan(ds["terminals"], ds.get("timeouts"), 60)
    np.testing.assert_array_equal(out["observations"], ds["observations"][:-1][keep])


@pytest.mark.parametrize("E,heads,I,L,q", [(64, 4, 128, 1, 25), (128, 2, 256, 3, 16), (256, 4, 1024, 1, 100)])
def test_pt_relabel_vs_oracle_larger_shapes(E, heads, I, L, q):
    """Left-padded and full windows, distinct windows (no dedupe) and the last-token-only final layer."""
    from iqlpref_b200.iql import RewardPT
    from iqlpref_b200.relabel import pt_predict
    S, A, n_data = 29, 8, 400
    torch.manual_seed(E + L)
    m = RewardPT(state_dim=S, action_dim=A, max_episode_steps=q + 50, embd_dim=E, pref_attn_embd_dim=E, num_heads=heads,
                 intermediate_dim=I, num_layers=L, max_pos=2 * q + 8, eps=0.1)
    with torch.no_grad():
        for n_, p_ in m.named_parameters():
            if "layer_norm" in n_ or n_.endswith("bias"):
                p_.add_(torch.randn_like(p_) * 0.2)
    rng = np.random.default_rng(E)
    obs = rng.standard_normal((n_data, S)).astype(np.float32)
    act = rng.uniform(-1, 1, (n_data, A)).astype(np.float32)
    n = 48
    seq = rng.integers(1, q + 1, n).astype(np.int32)
    seq[:4] = [1, q, q, 2]
    start = np.array([rng.integers(0, n_data - s + 1) for s in seq], dtype=np.int32)
    got = pt_predict(m.cuda().eval(), obs, act, start, seq, q, dedupe=False).cpu().numpy()
    # oracle: build the explicit windows the way iql.py:1264-1292 does
    sts = np.zeros((n, q, S), np.float32); acts = np.zeros((n, q, A), np.float32)
    ts = np.zeros((n, q), np.int64); am = np.zeros((n, q), np.float32)
    for j in range(n):
        pad = q - seq[j]
        sts[j, pad:] = obs[start[j]:start[j] + seq[j]]; acts[j, pad:] = act[start[j]:start[j] + seq[j]]
        ts[j, pad:] = np.arange(seq[j]); am[j, pad:] = 1
    p = {k: v.detach().cpu().numpy() for k, v in m.state_dict().items()}
    want = ro.pt_forward_value(p, sts, acts, ts, am, heads, 0.1)[:, -1]
    np.testing.assert_allclose(got, want, rtol=1e-3, atol=2e-5)
    # dedupe path: duplicated windows map back to the same values
    dup = np.concatenate([np.arange(n), np.arange(n)[::-1]])
    got2 = pt_predict(m, obs, act, start[dup], seq[dup], q).cpu().numpy()
    np.testing.assert_allclose(got2, got[dup], rtol=0, atol=0)
    with pytest.raises(ValueError):
        pt_predict(m, obs, act, np.array([n_data - 1], np.int32), np.array([2], np.int32), q)
