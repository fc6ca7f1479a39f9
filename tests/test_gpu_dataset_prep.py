"""GPU parity of the dataset preparation (return_reward_range / modify_reward / compute_mean_std / normalize_states,
iql.py:132-139, 344-401).  Episode returns are summed in double in the reference's order, so min / max returns and the
episode lengths are compared for equality; modify_reward replays numpy's op sequence and dtypes and must reproduce the
reference fixtures (tests/golden/relabel.npz, written by the unmodified reference) bit for bit."""
import os

import numpy as np
import pytest
import torch

from oracle import relabel_oracle as ro

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def rel(golden_dir):
    return np.load(os.path.join(golden_dir, "relabel.npz"))


@pytest.mark.parametrize("env", ["antmaze-large-diverse-v2", "halfcheetah-medium-v2"])
@pytest.mark.parametrize("nr", list(range(0, 9)))
def test_modify_reward_reproduces_the_reference(rel, env, nr):
    from iqlpref_b200 import iql as biql
    d = {"rewards": rel["ds_t.rewards"].copy(), "terminals": rel["ds_t.terminals"].copy()}
    biql.modify_reward(d, env, nr, max_episode_steps=50)
    want = rel[f"modrew.{env}.{nr}"] if nr >= 1 else ro.modify_reward(rel["ds_t.rewards"], rel["ds_t.terminals"], env, 0, 50)
    assert d["rewards"].dtype == want.dtype and d["rewards"].shape == want.shape
    np.testing.assert_array_equal(d["rewards"], want)


def test_modify_reward_leaves_other_environments_alone(rel):
    from iqlpref_b200 import iql as biql
    d = {"rewards": rel["ds_t.rewards"].copy(), "terminals": rel["ds_t.terminals"].copy()}
    biql.modify_reward(d, "pen-human-v1", 3, max_episode_steps=50)
    np.testing.assert_array_equal(d["rewards"], rel["ds_t.rewards"])


@pytest.mark.parametrize("n,p_term,cap", [(200_000, 0.002, 1000), (50_000, 0.0, 700), (30_001, 0.05, 50), (1500, 0.01, 1),
                                          (5, 0.0, 1000)])
def test_return_reward_range_equals_the_sequential_loop(n, p_term, cap):
    from iqlpref_b200 import iql as biql
    rng = np.random.default_rng(n)
    rewards = (rng.standard_normal(n) * 3).astype(np.float32)
    terminals = rng.uniform(size=n) < p_term
    if n == 5:  # nothing ever completes: the reference's min([]) raises
        with pytest.raises(ValueError):
            biql.return_reward_range({"rewards": rewards, "terminals": terminals}, cap)
        return
    mn, mx, trj = biql.return_reward_range({"rewards": rewards, "terminals": terminals}, cap)
    w_mn, w_mx, w_trj = ro.return_reward_range(rewards, terminals, cap)
    assert mn == w_mn and mx == w_mx          # same additions in the same order, in double
    np.testing.assert_array_equal(trj, w_trj)


def test_reward_range_full_size_properties():
    """1M rows (BASELINE configs[1] size): with no terminals every episode is exactly `cap` rows, and shifting all
    rewards by a constant shifts every return by cap * constant."""
    from iqlpref_b200.dataset_prep import return_reward_range_device
    n, cap = 1_000_000, 1000
    g = torch.Generator(device="cuda").manual_seed(0)
    r = torch.randn(n, generator=g, device="cuda")
    t = torch.zeros(n, dtype=torch.uint8, device="cuda")
    mm, trj = return_reward_range_device(r, t, cap)
    assert torch.all(trj == cap)
    ret = r.double().view(-1, cap).sum(1)
    torch.testing.assert_close(mm, torch.stack([ret.min(), ret.max()]), rtol=1e-12, atol=1e-9)
    mm2, _ = return_reward_range_device(r + 0.5, t, cap, want_lengths=False)
    torch.testing.assert_close(mm2, mm + 0.5 * cap, rtol=1e-6, atol=1e-3)


def test_mean_std_and_normalize_match_numpy():
    from iqlpref_b200 import iql as biql
    from iqlpref_b200.dataset_prep import compute_mean_std_device, normalize_states_device
    rng = np.random.default_rng(1)
    for n, dim in ((100_000, 29), (3, 45), (1, 17)):
        x = (rng.standard_normal((n, dim)) * rng.uniform(0.1, 5, dim) + rng.uniform(-3, 3, dim)).astype(np.float32)
        mean, std = biql.compute_mean_std(x, 1e-3)
        x64 = x.astype(np.float64)
        np.testing.assert_allclose(mean, x64.mean(0), rtol=1e-6, atol=1e-6)   # accumulated in double: closer to the
        np.testing.assert_allclose(std, x64.std(0) + 1e-3, rtol=1e-6, atol=1e-6)  # exact value than numpy's fp32 pass
        np.testing.assert_allclose(mean, x.mean(0), rtol=1e-4, atol=1e-4)
        got = biql.normalize_states(x, mean, std)
        np.testing.assert_array_equal(got, (x - mean) / std)  # elementwise fp32: bit-exact
    # in place on a row-strided view of a packed replay table
    table = torch.randn((5000, 68), device="cuda")
    view = table[:, :29]
    before = view.clone()
    m, s = compute_mean_std_device(view, 1e-3)
    normalize_states_device(view, m, s)
    torch.testing.assert_close(view, (before - m) / s, rtol=0, atol=0)
    torch.testing.assert_close(table[:, 29:], table[:, 29:])
