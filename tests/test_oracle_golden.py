"""Pin the CPU oracle (oracle/) against fixtures produced by the UNMODIFIED reference
(oracle/gen_golden.py, run in the authoring container).  CPU only."""
import os

import numpy as np
import pytest
import torch
import yaml

from oracle import relabel_oracle as ro
from oracle.iql_oracle import IQLOracle, OracleHParams
from oracle.philox_oracle import cuda_randint_indices, offset_increment, philox4x32_10


def test_philox_known_answers():
    # Random123 kat_vectors for philox4x32-10
    kat = [
        ([0, 0, 0, 0], [0, 0], [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]),
        ([0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2, [0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD]),
        ([0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344], [0xA4093822, 0x299F31D0],
         [0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1]),
    ]
    for c, k, want in kat:
        got = philox4x32_10(np.array([c], dtype=np.uint64), np.array([k], dtype=np.uint64))[0]
        assert [int(x) for x in got] == want


def test_randint_stream_shape_and_range():
    idx = cuda_randint_indices(seed=5, offset=8, n=999_983, batch=256)
    assert idx.dtype == np.int64 and idx.shape == (256,)
    assert idx.min() >= 0 and idx.max() < 999_983
    assert offset_increment(256, 1_000_000) == 4
    # 64-bit path for huge ranges
    big = cuda_randint_indices(seed=5, offset=8, n=(1 << 28) + 7, batch=16)
    assert big.max() < (1 << 28) + 7


def _nets_from_flat(d, prefix, S, A, det):
    def net(root, idxs):
        return [(torch.from_numpy(d[f"{prefix}.{root}.{i}.weight"]), torch.from_numpy(d[f"{prefix}.{root}.{i}.bias"]))
                for i in idxs]
    q1 = net("qf.q1.net", (0, 2, 4))
    q2 = net("qf.q2.net", (0, 2, 4))
    v = net("vf.v.net", (0, 2, 4))
    a = net("actor.net.net", (0, 2, 4))
    ls = None if det else torch.from_numpy(d[f"{prefix}.actor.log_std"])
    return q1, q2, v, a, ls


@pytest.mark.parametrize("name,det", [("gauss", False), ("det", True)])
def test_iql_oracle_matches_reference_small(golden_dir, name, det):
    d = np.load(os.path.join(golden_dir, f"iql_small_{name}_fp32.npz"))
    hp = yaml.safe_load(str(d["hp"]))
    orc = IQLOracle(*_nets_from_flat(d, "init", 17, 6, det), OracleHParams(deterministic=det, **hp), mode="fp32")
    tab = {k: torch.from_numpy(d[f"table.{k}"]) for k in ("obs", "act", "rew", "next_obs", "done")}
    for k in range(d["idx"].shape[0]):
        i = torch.from_numpy(d["idx"][k])
        out = orc.step(tab["obs"][i], tab["act"][i], tab["rew"][i], tab["next_obs"][i], tab["done"][i])
        np.testing.assert_allclose([out["value_loss"], out["q_loss"], out["actor_loss"]], d["losses"][k], rtol=1e-6)
    e = orc.export()
    np.testing.assert_allclose(e["q1.w1"], d["final.qf.q1.net.2.weight"], rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(e["qt2.w0"], d["final.q_target.q2.net.0.weight"], rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(e["v.b2"], d["final.vf.v.net.4.bias"], rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(e["actor.w2"], d["final.actor.net.net.4.weight"], rtol=1e-5, atol=1e-7)
    if not det:
        np.testing.assert_allclose(e["log_std"], d["final.actor.log_std"], rtol=1e-5, atol=1e-7)
    assert abs(orc.a_opt.lr - float(d["lr_final"])) < 1e-15


@pytest.fixture(scope="module")
def rel(golden_dir):
    return np.load(os.path.join(golden_dir, "relabel.npz"))


@pytest.mark.parametrize("act", ["relu", "tanh", "swish", "softplus", "leaky_relu", "rbf", "sin", "cos", "linear"])
def test_reward_mlp_forward(rel, act):
    w = [rel[f"mlp.{act}.w{i}"] for i in range(8)]
    y = ro.reward_mlp_forward(rel["mlp.x"], w, act)
    np.testing.assert_allclose(y, rel[f"mlp.{act}.y"], rtol=2e-5, atol=2e-6)


def _ds(rel, tag):
    d = {k: rel[f"{tag}.{k}"] for k in ("observations", "actions", "rewards", "terminals")}
    d["timeouts"] = rel[f"{tag}.timeouts"] if f"{tag}.timeouts" in rel.files else None
    return d


@pytest.mark.parametrize("tag", ["ds_t", "ds_n"])
@pytest.mark.parametrize("toe", [False, True])
def test_keep_mask_and_mr_relabel(rel, tag, toe):
    ds = _ds(rel, tag)
    keep, _ = ro.keep_mask_scan(ds["terminals"], ds["timeouts"], 60, terminate_on_end=toe)
    want_obs = rel[f"mr.{tag}.toe{int(toe)}.observations"]
    np.testing.assert_array_equal(ds["observations"][:-1][keep], want_obs)
    np.testing.assert_array_equal(ds["observations"][1:][keep], rel[f"mr.{tag}.toe{int(toe)}.next_observations"])
    x = np.concatenate([ds["observations"][:-1], ds["actions"][:-1]], 1)
    r = ro.reward_mlp_forward(x, [rel[f"mr.w{i}"] for i in range(8)], "relu")[:, 0]
    np.testing.assert_allclose(r[keep], rel[f"mr.{tag}.toe{int(toe)}.rewards"], rtol=2e-5, atol=2e-6)


@pytest.mark.parametrize("alpha", [0.0, 0.5, 0.9, 0.95])
@pytest.mark.parametrize("burn", [0, 10])
def test_mr_ensemble_cvar(rel, alpha, burn):
    ds = _ds(rel, "ds_t")
    keep, _ = ro.keep_mask_scan(ds["terminals"], ds["timeouts"], 60)
    x = np.concatenate([ds["observations"][:-1], ds["actions"][:-1]], 1)
    members = [[rel[f"ens.e{e}.w{i}"] for i in range(6)] for e in range(0, 60, 5) if e >= burn]
    r = ro.ensemble_relabel(x, members, "tanh", alpha)
    np.testing.assert_allclose(r[keep], rel[f"ens.a{alpha}.b{burn}.rewards"], rtol=2e-5, atol=2e-6)


@pytest.mark.parametrize("alpha,ns", [(0.9, 500), (0.9, 16), (0.0, 40), (0.75, 40)])
def test_bnn_cvar(rel, alpha, ns):
    ds = _ds(rel, "ds_t")
    keep, _ = ro.keep_mask_scan(ds["terminals"], ds["timeouts"], 60)
    x = np.concatenate([ds["observations"][:-1], ds["actions"][:-1]], 1)
    members = [[rel[f"bnn.c{c}.k{k}.w{i}"] for i in range(8)] for c in range(2) for k in range(20)]
    if 0 < ns < len(members):  # iql.py:929-932
        idx = np.random.default_rng(seed=0).choice(len(members), size=ns, replace=False)
        members = [members[i] for i in sorted(idx)]
    r = ro.ensemble_relabel(x, members, "relu", alpha)
    np.testing.assert_allclose(r[keep], rel[f"bnn.a{alpha}.n{ns}.rewards"], rtol=2e-5, atol=2e-6)


def _pt_params(rel, L):
    pre = f"pt{L}."
    return {k[len(pre):]: rel[k] for k in rel.files if k.startswith(pre) and ".fw." not in k and ".ds_" not in k}


@pytest.mark.parametrize("L", [1, 2])
def test_pt_forward_value(rel, L):
    p = _pt_params(rel, L)
    v = ro.pt_forward_value(p, rel[f"pt{L}.fw.sts"], rel[f"pt{L}.fw.acts"], rel[f"pt{L}.fw.ts"], rel[f"pt{L}.fw.am"], 4, 0.1)
    np.testing.assert_allclose(v, rel[f"pt{L}.fw.value"], rtol=1e-4, atol=1e-5)
    v2, w = ro.pt_forward_value(p, rel[f"pt{L}.fw.sts"], rel[f"pt{L}.fw.acts"], rel[f"pt{L}.fw.ts"], rel[f"pt{L}.fw.am"], 4,
                                0.1, with_weighted_sum=True)
    np.testing.assert_array_equal(v2, v)
    np.testing.assert_allclose(w, rel[f"pt{L}.fw.weighted_sum"], rtol=1e-4, atol=1e-5)  # RewardPT's preference attention


@pytest.mark.parametrize("L", [1, 2])
@pytest.mark.parametrize("tag", ["ds_t", "ds_n"])
def test_pt_relabel(rel, L, tag):
    ds = _ds(rel, tag)
    keep, ep_steps = ro.keep_mask_scan(ds["terminals"], ds["timeouts"], 60)
    r = ro.pt_relabel(_pt_params(rel, L), ds["observations"], ds["actions"], ep_steps, 10, 4, 0.1)
    np.testing.assert_allclose(r[keep], rel[f"pt{L}.{tag}.rewards"], rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("env", ["antmaze-large-diverse-v2", "halfcheetah-medium-v2"])
@pytest.mark.parametrize("nr", list(range(1, 9)))
def test_modify_reward(rel, env, nr):
    ds = _ds(rel, "ds_t")
    r = ro.modify_reward(ds["rewards"], ds["terminals"], env, nr, max_episode_steps=50)
    np.testing.assert_allclose(r, rel[f"modrew.{env}.{nr}"], rtol=1e-6, atol=1e-7)


def test_empirical_cvar(rel):
    s = rel["cvar.samples"]
    got = [ro.cvar_reduce(s[:, None], a)[0] for a in (0.0, 0.5, 0.9, 0.95, 0.99)]
    np.testing.assert_allclose(got, rel["cvar.values"], rtol=1e-6)


# ---- checkpoint format (iql.py:664-688): the oracle resumes from the reference's own file ---------------------
def test_oracle_resumes_from_the_reference_checkpoint(golden_dir):
    import yaml
    from oracle.iql_oracle import IQLOracle, OracleHParams
    d = np.load(os.path.join(golden_dir, "iql_small_ckpt_resume.npz"))
    sd = torch.load(os.path.join(golden_dir, "iql_small_ckpt.pt"), map_location="cpu", weights_only=False)
    hp = yaml.safe_load(str(d["hp"]))
    S, A, H = 17, 6, 64
    z = lambda o, i: (torch.zeros(o, i), torch.zeros(o))
    net = lambda i, o: [z(H, i), z(H, H), z(o, H)]
    orc = IQLOracle(net(S + A, 1), net(S + A, 1), net(S, 1), net(S, A), torch.zeros(A), OracleHParams(**hp), mode="fp32")
    orc.load_checkpoint(sd)
    assert orc.total_it == int(d["k_saved"]) and orc.a_opt.step_count == int(d["k_saved"])
    t = {n: torch.from_numpy(d[f"table.{n}"]) for n in ("obs", "act", "rew", "next_obs", "done")}
    for k in range(int(d["k_saved"]), d["idx"].shape[0]):
        i = torch.from_numpy(d["idx"][k])
        out = orc.step(t["obs"][i], t["act"][i], t["rew"][i], t["next_obs"][i], t["done"][i])
        np.testing.assert_allclose([out["value_loss"], out["q_loss"], out["actor_loss"]], d["losses"][k], rtol=1e-5)
    e = orc.export()
    np.testing.assert_allclose(e["q1.w1"], d["final.qf.q1.net.2.weight"], rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(e["qt2.w0"], d["final.q_target.q2.net.0.weight"], rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(e["log_std"], d["final.actor.log_std"], rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(orc.a_opt.exp_avg[0].numpy(), d["final.actor_optimizer.0.exp_avg"], rtol=1e-4, atol=1e-9)
    assert abs(orc.a_opt.lr - float(d["lr_final"])) < 1e-15


def test_mirror_modules_enumerate_parameters_like_the_reference_checkpoint(golden_dir):
    """The optimizer state of a checkpoint is indexed by position in `parameters()`: the host mirror's modules must
    enumerate names in the order the reference's file holds them (`log_std` first for the Gaussian policy)."""
    from iqlpref_b200 import iql as biql
    sd = torch.load(os.path.join(golden_dir, "iql_small_ckpt.pt"), map_location="cpu", weights_only=False)
    mods = {"qf": biql.TwinQ(17, 6, 64), "vf": biql.ValueFunction(17, 64), "actor": biql.GaussianPolicy(17, 6, 1.0, 64)}
    for part, opt in (("qf", "q_optimizer"), ("vf", "v_optimizer"), ("actor", "actor_optimizer")):
        names = [n for n, _ in mods[part].named_parameters()]
        assert names == list(sd[part])
        for i, (n, prm) in enumerate(mods[part].named_parameters()):
            assert tuple(sd[opt]["state"][i]["exp_avg"].shape) == tuple(prm.shape) == tuple(sd[part][n].shape)
