"""The MRTrainerF oracle (oracle/mr_train_oracle.py: MRTrainerFOracle, woodbury_solve, lcf_log_prior) against
tests/golden/mrf_train.npz, written by the unmodified reference MRTrainerF + LCFModel (oracle/gen_golden_mrf.py)."""
import os

import numpy as np
import pytest

from oracle.gen_golden_mrf import CASES, load_case
from oracle.mr_train_oracle import MRTrainerFOracle, synth_features, woodbury_solve

GOLD = os.path.join(os.path.dirname(__file__), "golden", "mrf_train.npz")


@pytest.mark.parametrize("name", list(CASES))
def test_oracle_reproduces_the_reference_mrtrainerf(name):
    c = load_case(np.load(GOLD), name)
    o = MRTrainerFOracle(c["w0"], c["act"], c["N"], c["p_mean"], c["sigma"], bt_pool=c["pool"])
    phi0 = synth_features(c["meas"][0], c["d"])
    _, _, g = o.grads(c["batches"][0], c["meas"][0], phi0)
    # the prior term is J^T alpha / N with alpha = r_perp / eps, eps ~ 1e-6: gradients are O(1e3-1e5) and fp32 forward noise
    # enters relatively.  Where the prior term cancels (the intercept column of Phi makes sum(alpha) ~ 0: output bias), what
    # is left is rounding noise of that scale, hence the second term.
    gmax = max(np.abs(w).max() for w in c["g0"])
    for i, (got, want) in enumerate(zip(g, c["g0"])):
        want = want.reshape(got.shape)
        assert np.abs(got - want).max() <= 2e-4 * np.abs(want).max() + 1e-6 * gmax, (i, np.abs(got - want).max(), np.abs(want).max())
    for k in range(c["K"]):
        log = o.train(c["batches"][k], c["meas"][k], synth_features(c["meas"][k], c["d"]))
        assert log["training_acc"] == pytest.approx(c["acc"][k], abs=1e-6)
        assert log["log_prior"] == pytest.approx(c["log_prior"][k], rel=2e-3)
        assert log["training_loss"] == pytest.approx(c["loss"][k], rel=2e-3)
    for i, (got, want) in enumerate(zip(o.weights, c["wK"])):
        # Adam normalises the gradient: after K steps every weight has moved by <= K * lr; allow 5 % of that
        assert np.abs(got - want.reshape(got.shape)).max() <= 0.05 * c["K"] * 3e-4, i
    ev = o.evaluation(c["batches"][c["K"]])
    assert ev["eval_loss"] == pytest.approx(c["eval"][0], rel=1e-3)


def test_woodbury_matches_the_dense_solve():
    rng = np.random.default_rng(0)
    phi = synth_features(rng.standard_normal((30, 8)), 6)
    sigma = np.diag(rng.uniform(0.5, 2, 6))
    r = rng.standard_normal(30)
    jitter = 1e-2  # a nugget large enough for the dense n x n system to be well conditioned
    alpha, logdet = woodbury_solve(phi, r, sigma, jitter)
    ptp = phi.T @ phi
    eps = jitter * max(np.einsum("ij,ij->", sigma, ptp) / 30, 1.0)
    K = phi @ sigma @ phi.T + eps * np.eye(30)
    np.testing.assert_allclose(alpha, np.linalg.solve(K, r), rtol=1e-8, atol=1e-10)
    assert logdet == pytest.approx(np.linalg.slogdet(K)[1], rel=1e-10)
    assert woodbury_solve(phi[:0], r[:0], sigma)[0].shape == (0,)
