"""CPU: the functional PT training oracle (oracle/pt_train_oracle.py) against fixtures written by the unmodified reference
`PTTrainer` on its own `PT` (tests/golden/pt_train.npz, oracle/gen_golden_pt.py): first-step gradients of every parameter,
per-step loss and accuracy, `evaluation()`, weights after K Adam steps."""
import os

import numpy as np
import pytest

from oracle.gen_golden_pt import CASES, load_case
from oracle.pt_train_oracle import PTTrainerOracle


def assert_weights_close(got, c, lr=3e-4):
    """Weights after K Adam steps: 3e-5 absolute, except elements whose gradient is rounding noise - Adam normalises
    any gradient to a step of ~lr, so those move by up to lr per step in either direction (bound 2 K lr).  Such elements
    exist by construction: a softmax is invariant to a constant added to all its scores, so the KEY biases of the causal
    attention and of the preference attention have gradient exactly 0 in real arithmetic, and with equal segment lengths
    so does the value bias under sum pooling.  They are recognised by the reference's own first-step gradient."""
    gmax = max(float(np.abs(g).max()) for g in c["g0"].values())
    for k, want in c["wK"].items():
        d = np.abs(np.asarray(got[k]) - want)
        assert d.max() <= 2 * c["K"] * lr, k
        real = np.abs(c["g0"][k]) > 1e-5 * gmax
        assert np.sum(d[real] > 3e-5) <= max(1, 0.02 * d.size), k


@pytest.fixture(scope="module")
def golden(golden_dir):
    return np.load(os.path.join(golden_dir, "pt_train.npz"))


@pytest.mark.parametrize("case", sorted(CASES))
def test_pt_train_oracle_matches_reference(golden, case):
    c = load_case(golden, case)
    orc = PTTrainerOracle(c["w0"], c["heads"], c["eps"], lr=3e-4, bt_pool=c["pool"])
    _, _, g = orc.loss_and_grads(c["batches"][0])
    assert set(g) == set(c["g0"])
    for k, want in c["g0"].items():
        np.testing.assert_allclose(g[k].numpy(), want, rtol=1e-4, atol=1e-6 * max(1.0, float(np.abs(want).max())), err_msg=k)
    for k in range(c["K"]):
        log = orc.train(c["batches"][k])
        assert log["training_loss"] == pytest.approx(c["loss"][k], rel=1e-5)
        assert log["training_acc"] == pytest.approx(c["acc"][k], abs=1e-7)
    ev = orc.evaluation(c["batches"][c["K"]])
    assert ev["eval_loss"] == pytest.approx(c["eval"][0], rel=1e-5) and ev["eval_acc"] == pytest.approx(c["eval"][1], abs=1e-7)
    assert_weights_close(orc.weights(), c)
