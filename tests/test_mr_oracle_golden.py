"""The MR-training oracle against fixtures written by the unmodified reference MRTrainer (CPU).

`tests/golden/mr_train.npz` comes from `oracle/gen_golden_mr.py` (reference `MRTrainer.train/.evaluation`,
gp_reward-priors/optbnn/training/training.py:27-114, torch Adam, CPU fp32)."""
import os

import numpy as np
import pytest

from oracle.gen_golden_mr import CASES, load_case
from oracle.mr_train_oracle import MRTrainerOracle, mr_loss_and_grads

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "mr_train.npz")


@pytest.fixture(scope="module")
def golden():
    return np.load(GOLDEN)


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_matches_reference_trainer(golden, name):
    c = load_case(golden, name)
    loss, acc, grads = mr_loss_and_grads(c["w0"], c["batches"][0], c["act"], c["pool"])
    assert abs(loss - c["loss"][0]) <= 1e-5 * max(1.0, abs(c["loss"][0]))
    assert acc == pytest.approx(c["acc"][0], abs=1e-6)
    for g, gr in zip(grads, c["g0"]):
        np.testing.assert_allclose(g.reshape(gr.shape), gr, rtol=1e-3, atol=1e-6 * max(1.0, float(np.abs(gr).max())))
    tr = MRTrainerOracle(c["w0"], c["act"], lr=3e-4, bt_pool=c["pool"])
    for k in range(c["K"]):
        log = tr.train(c["batches"][k])
        assert log["training_loss"] == pytest.approx(c["loss"][k], rel=1e-4)
        assert log["training_acc"] == pytest.approx(c["acc"][k], abs=1e-6)
    ev = tr.evaluation(c["batches"][c["K"]])
    assert ev["eval_loss"] == pytest.approx(c["eval"][0], rel=1e-4)
    assert ev["eval_acc"] == pytest.approx(c["eval"][1], abs=1e-6)
    for w, wr, m, mr, v, vr in zip(tr.weights, c["wK"], tr.m, c["mK"], tr.v, c["vK"]):
        # Adam moves an element by ~lr per step whatever the gradient's size, so an element whose gradient is
        # rounding noise may differ by up to K*lr; everything else agrees to 3e-5
        d = np.abs(w - wr)
        assert d.max() <= 2 * c["K"] * 3e-4
        # (the output bias is the extreme case: its Bradley-Terry gradient is exactly 0 in real arithmetic)
        assert np.sum(d > 3e-5) <= max(1, 0.02 * d.size)
        np.testing.assert_allclose(m, mr, rtol=1e-3, atol=1e-6 * max(1.0, float(np.abs(mr).max())))
        np.testing.assert_allclose(v, vr, rtol=2e-3, atol=1e-9)


def test_pool_mode_error():
    from oracle.mr_train_oracle import mr_forward, synth_batch
    b = synth_batch(np.random.default_rng(0), 2, 3, 2, 1)
    w = [np.zeros((3, 4), np.float32), np.zeros(4, np.float32), np.zeros((4, 1), np.float32), np.zeros(1, np.float32)]
    with pytest.raises(ValueError):
        mr_forward(w, b, "relu", "median")
