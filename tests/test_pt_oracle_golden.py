"""CPU: the functional PT training oracle (oracle/pt_train_oracle.py) against fixtures written by the unmodified reference
`PTTrainer` on its own `PT` (tests/golden/pt_train.npz, oracle/gen_golden_pt.py): first-step gradients of every parameter,
per-step loss and accuracy, `evaluation()`, weights after K Adam steps."""
import os

import numpy as np
import pytest

from oracle.gen_golden_pt import CASES, load_case
from oracle.pt_train_oracle import PTTrainerOracle


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
    for k, want in c["wK"].items():  # (Adam moves an element whose gradient is rounding noise by ~lr per step either way)
        d = np.abs(orc.weights()[k] - want)
        assert d.max() <= 2 * c["K"] * 3e-4 and np.sum(d > 3e-5) <= max(1, 0.02 * d.size), k
