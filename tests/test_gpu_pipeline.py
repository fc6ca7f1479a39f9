"""BASELINE configs[2] and configs[3] end to end at toy size (scripts/pipeline_cfg3_cfg4.py): preference-transformer training ->
best_model.pt -> load_pt_reward_model -> qlearning_dataset_pt -> IQL, and MR snapshots -> qlearning_dataset_mr_ensemble (CVaR)
-> IQL with actor dropout.  Every stage is parity-tested on its own elsewhere; this checks that they compose (file formats,
device placement, shapes) and that the losses come out finite."""
import json
import os
import sys

import pytest

pytestmark = pytest.mark.gpu


def test_pipelines_compose(monkeypatch, capsys):
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    monkeypatch.syspath_prepend(os.path.join(root, "scripts"))
    monkeypatch.setattr(sys, "argv", ["pipeline", "--rows", "6000", "--agents", "2", "--iql-steps", "5", "--epochs", "1", "--batches", "2"])
    import pipeline_cfg3_cfg4 as pipe
    with pytest.warns(RuntimeWarning):  # 10 snapshots at alpha 0.9: the reference's "CVaR tail has only 1 snapshot" warning
        pipe.main()
    lines = [json.loads(l) for l in capsys.readouterr().out.splitlines() if l.startswith("{")]
    assert len(lines) == 2 and "configs[2]" in lines[0]["pipeline"] and "configs[3]" in lines[1]["pipeline"]
    for l in lines:
        assert all(abs(x) < 1e6 for x in l["final_losses_agent0"])
    assert lines[0]["reward_range"][0] < lines[0]["reward_range"][1]
