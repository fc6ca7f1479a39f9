"""GPU parity of `MRTrainerF` (training.py:132-305 with the LCF prior of gp/models/model.py:202-513):
`iqlpref_b200.mr_training.MRTrainerF` (`iqlb_mr_train_step_f`) against fixtures written by the unmodified reference
MRTrainerF + LCFModel (tests/golden/mrf_train.npz, oracle/gen_golden_mrf.py) and against the oracle on fresh inputs
(more concepts, a production-sized batch, split-bf16 GEMMs).

Tolerances.  The prior term of the gradient is J^T alpha / N with alpha = r_perp / eps, eps ~ 1e-6, so gradients are
O(1e3-1e5) and the fp32 forward noise of f(X_M) enters relatively: 1e-3 of a tensor's largest entry, floored at 1e-5 of
the whole gradient's largest entry (where the prior term cancels - the intercept column of Phi makes sum(alpha) ~ 0, the
output-bias gradient - what is left is rounding noise of that scale).  log p_GP and the logged energy: 2e-3 relative.
Weights after K Adam steps: Adam normalises the gradient, every weight moves by <= lr per step; 3e-5 absolute except
noise-gradient elements (<= 2 %, bounded by 2*K*lr), the bar of tests/test_gpu_mr_training.py."""
import os

import numpy as np
import pytest
import torch

from oracle.gen_golden_mrf import CASES, load_case
from oracle.mr_train_oracle import MRTrainerFOracle, synth_batch, synth_features

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def golden(golden_dir):
    return np.load(os.path.join(golden_dir, "mrf_train.npz"))


def feature_fn(d):
    def fv(X, device):  # the reference signature: function_vect(X, device) -> (n, d)
        return torch.from_numpy(synth_features(X.detach().cpu().numpy(), d)).to(device)
    return fv


def make_trainer(weights, act, pool, N, p_mean, sigma, d, precision="fp32", wrap_reference_like=False):
    from iqlpref_b200.gp_prior import LCFModel
    from iqlpref_b200.mr_training import MRTrainerF
    from iqlpref_b200.reward_models.q_mlp import MLP
    d_in = weights[0].shape[0]
    widths = [w.shape[1] for w in weights[0:-2:2]]
    net = MLP(d_in, 1, widths, act)
    with torch.no_grad():
        for p, w in zip(net.parameters(), weights):
            p.copy_(torch.from_numpy(np.asarray(w)).reshape(p.shape))
    opt = torch.optim.Adam(net.parameters(), lr=3e-4)
    if wrap_reference_like:  # an object shaped like the reference LCFModel (tensors + callable) is re-wrapped
        class Prior:
            pass
        prior = Prior()
        prior.p_mean, prior.p_covariance, prior.function_vect = torch.from_numpy(p_mean), torch.from_numpy(sigma), feature_fn(d)
    else:
        prior = LCFModel(sigma, feature_fn(d), "cuda", p_mean=p_mean)
    return MRTrainerF(net, opt, N, prior, device="cuda", bt_pool=pool, precision=precision), net, opt


def tb(batch):
    return [torch.from_numpy(np.asarray(x)) for x in batch]


def split(flat, weights):
    out, off = [], 0
    flat = flat.detach().cpu().numpy()
    for w in weights:
        out.append(flat[off:off + w.size].reshape(w.shape)); off += w.size
    return out


def assert_grads_close(got, want, rel=1e-3, floor_rel=1e-5):
    gmax = max(float(np.abs(np.asarray(w)).max()) for w in want)
    for i, (g, w) in enumerate(zip(got, want)):
        w = np.asarray(w).reshape(g.shape)
        tol = rel * float(np.abs(w).max()) + floor_rel * gmax
        assert np.abs(g - w).max() <= tol, (i, float(np.abs(g - w).max()), tol)


def assert_params_close(got, want, K, lr=3e-4, frac=0.02):
    for w, wr in zip(got, want):
        d = np.abs(np.asarray(w).reshape(-1) - np.asarray(wr).reshape(-1))
        assert d.max() <= 2 * K * lr + 1e-6
        assert np.sum(d > 3e-5) <= max(1, frac * d.size), (float(d.max()), int(np.sum(d > 3e-5)), d.size)


@pytest.mark.parametrize("name", sorted(CASES))
def test_reference_fixture(golden, name):
    c = load_case(golden, name)
    tr, net, opt = make_trainer(c["w0"], c["act"], c["pool"], c["N"], c["p_mean"], c["sigma"], c["d"],
                                wrap_reference_like=(name == "tanh_full"))
    for k in range(c["K"]):
        log = tr.train(tb(c["batches"][k]), c["meas"][k], return_grad=True)
        assert log["training_acc"] == pytest.approx(c["acc"][k], abs=1e-6)
        assert log["log_prior"] == pytest.approx(c["log_prior"][k], rel=2e-3)
        assert log["training_loss"] == pytest.approx(c["loss"][k], rel=2e-3)
        if k == 0:
            assert_grads_close(split(log["grad"], c["w0"]), c["g0"])
    ev = tr.evaluation(tb(c["batches"][c["K"]]))
    assert ev["eval_loss"] == pytest.approx(c["eval"][0], rel=1e-3)
    assert ev["eval_acc"] == pytest.approx(c["eval"][1], abs=1e-6)
    assert_params_close([p.detach().cpu().numpy() for p in net.parameters()], c["wK"], c["K"])
    st = tr.state_dict()["optimizer"]["state"]
    assert float(st[0]["step"]) == c["K"]


@pytest.mark.parametrize("precision,B,T,n_meas,d", [("fp32", 24, 10, 100, 6), ("fp32", 5, 4, 7, 3),
                                                    ("split_bf16", 64, 25, 256, 6)])
def test_oracle_fresh_inputs(precision, B, T, n_meas, d):
    rng = np.random.default_rng(B * 1000 + n_meas)
    S, A, K = 9, 4, 3
    dims = [S + A, 128, 128, 1]
    weights = []
    for i in range(len(dims) - 1):
        weights += [rng.standard_normal((dims[i], dims[i + 1])).astype(np.float32),
                    (0.1 * rng.standard_normal(dims[i + 1])).astype(np.float32)]
    sigma = np.diag(rng.uniform(0.5, 2.0, d)).astype(np.float32)
    p_mean = (0.2 * rng.standard_normal(d)).astype(np.float32)
    N = 500
    o = MRTrainerFOracle(weights, "relu", N, p_mean, sigma)
    tr, net, _ = make_trainer(weights, "relu", "mean", N, p_mean, sigma, d, precision=precision)
    loose = precision == "split_bf16"
    for k in range(K):
        batch = synth_batch(rng, B, T, S, A)
        xm = rng.standard_normal((n_meas, S + A)).astype(np.float32)
        phi = synth_features(xm, d)
        if k == 0:
            _, _, g = o.grads(batch, xm, phi)
        want = o.train(batch, xm, phi)
        log = tr.train(tb(batch), xm, return_grad=True)
        if k == 0:
            assert_grads_close(split(log["grad"], weights), g, rel=1e-2 if loose else 1e-3, floor_rel=1e-4 if loose else 1e-5)
        assert log["training_acc"] == pytest.approx(want["training_acc"], abs=1e-6)
        assert log["ce_loss"] == pytest.approx(want["ce_loss"], rel=1e-3 if loose else 1e-4)
        assert log["log_prior"] == pytest.approx(want["log_prior"], rel=2e-2 if loose else 2e-3)
    assert_params_close([p.detach().cpu().numpy() for p in net.parameters()], o.weights, K, frac=0.05 if loose else 0.02)


def test_errors_and_async_log():
    rng = np.random.default_rng(5)
    S, A, d = 6, 2, 4
    weights = [rng.standard_normal((8, 32)).astype(np.float32), np.zeros(32, np.float32),
               rng.standard_normal((32, 1)).astype(np.float32), np.zeros(1, np.float32)]
    sigma, p_mean = np.ones(d, np.float32), np.zeros(d, np.float32)  # a (d,) covariance is a diagonal (model.py:236-241)
    tr, _, _ = make_trainer(weights, "tanh", "mean", 100, p_mean, sigma, d)
    batch = synth_batch(rng, 4, 5, S, A)
    xm = rng.standard_normal((9, S + A)).astype(np.float32)
    log = tr.train(tb(batch), xm, sync=False)
    assert torch.is_tensor(log["training_loss"]) and log["training_loss"].is_cuda and torch.isfinite(log["training_loss"])
    with pytest.raises(RuntimeError, match="cannot be multiplied"):
        tr.train(tb(batch), xm[:, :5])
    with pytest.raises(ValueError, match="empty measurement set"):
        tr.train(tb(batch), xm[:0])
    from iqlpref_b200.mr_training import MRTrainerF
    with pytest.raises(ValueError, match="prior is required"):
        MRTrainerF(tr.net, tr.opt, 1, None)
    # more concepts than the kernel's d x d shared-memory system holds: refused by the library
    from iqlpref_b200.gp_prior import LCFModel
    big = LCFModel(np.ones(65, np.float32), lambda X, device: torch.ones((X.shape[0], 65), dtype=torch.float64, device=device))
    tr.prior = big
    with pytest.raises(ValueError, match="n_concepts"):
        tr.train(tb(batch), xm)
