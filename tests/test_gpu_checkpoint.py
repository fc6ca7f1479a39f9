"""GPU parity of the checkpoint format: the file the reference's training loop writes
(`torch.save(trainer.state_dict(), checkpoint_{t}.pt)`, iql.py:1557-1561, :664-674) and what the reference does
after `load_state_dict` of it (iql.py:676-688).  Fixtures: tests/golden/iql_small_ckpt.pt (written by the
unmodified reference) and iql_small_ckpt_resume.npz (oracle/gen_golden_ckpt.py).  Tolerance 1e-3 (fp32 path)."""
import os

import numpy as np
import pytest
import torch
import yaml

pytestmark = pytest.mark.gpu

S, A, H, B = 17, 6, 64, 64
PARTS = (("qf", "q_optimizer"), ("vf", "v_optimizer"), ("actor", "actor_optimizer"))


def _close(got, want, rtol=1e-3, what=""):
    got, want = np.asarray(got, dtype=np.float64), np.asarray(want, dtype=np.float64)
    atol = rtol * max(float(np.abs(want).max()), 1e-6)  # at 1e-3 of the tensor's largest magnitude (cf. test_gpu_iql_step)
    err = np.abs(got - want)
    assert got.shape == want.shape, f"{what}: shape {got.shape} != {want.shape}"
    assert (err <= atol + rtol * np.abs(want)).all(), f"{what}: max err {err.max():.3e} (max |want| {np.abs(want).max():.3e})"


def _fixtures(golden_dir):
    d = np.load(os.path.join(golden_dir, "iql_small_ckpt_resume.npz"))
    sd = torch.load(os.path.join(golden_dir, "iql_small_ckpt.pt"), map_location="cpu", weights_only=False)
    return d, sd, yaml.safe_load(str(d["hp"]))


def _rows(biql, d, k, R):
    i = torch.from_numpy(d["idx"][k])
    t = {n: torch.from_numpy(d[f"table.{n}"]) for n in ("obs", "act", "rew", "next_obs", "done")}
    return biql.pack_rows(t["obs"][i], t["act"][i], t["rew"][i], t["next_obs"][i], t["done"][i], R).cuda()


def _modules(biql, d=None, seed=0):
    torch.manual_seed(seed)
    qf, vf, actor = biql.TwinQ(S, A, H), biql.ValueFunction(S, H), biql.GaussianPolicy(S, A, 1.0, H)
    if d is not None:
        for nm, mod in (("qf", qf), ("vf", vf), ("actor", actor)):
            mod.load_state_dict({k[len("init." + nm) + 1:]: torch.from_numpy(d[k]) for k in d.files
                                 if k.startswith(f"init.{nm}.")})
    return qf, vf, actor


def test_resume_from_a_reference_checkpoint(golden_dir):
    """Slot 1 of a two-agent engine loads the reference's file and continues like the reference does."""
    from iqlpref_b200 import iql as biql
    d, sd, hp = _fixtures(golden_dir)
    k0 = int(d["k_saved"])
    eng = biql.BatchedIQL(S, A, 2, hidden_dim=H, batch_size=B, hparams=[hp, hp])
    for a in range(2):
        eng.adopt(a, *_modules(biql, seed=40 + a), bind=False)
    assert eng.load_agent_state_dict(1, sd) == k0
    assert int(eng.it[1].item()) == k0 and int(eng.it[0].item()) == 0
    assert abs(float(eng.actor_lr[1].item()) - float(d["lr_saved"])) < 1e-15
    e = eng.export(1)
    np.testing.assert_array_equal(e["q1.w1"], sd["qf"]["q1.net.2.weight"].numpy())
    np.testing.assert_array_equal(e["qt1.w1"], e["q1.w1"])  # iql.py:679: the target restarts from qf
    np.testing.assert_array_equal(eng.view(eng.exp_avg, 1, "log_std").cpu().numpy(),
                                  sd["actor_optimizer"]["state"][0]["exp_avg"].numpy())  # log_std is parameter 0
    for k in range(k0, d["idx"].shape[0]):
        rows = _rows(biql, d, k, eng.R)
        eng.train(torch.stack([rows, rows]))
        _close(eng.losses[1].cpu().numpy(), d["losses"][k], what=f"losses of update {k}")
    out = eng.agent_state_dict(1)
    for part, opt in PARTS:
        for i, (name, val) in enumerate(out[part].items()):
            _close(val.cpu().numpy(), d[f"final.{part}.{name}"], what=f"{part}.{name}")
            st = out[opt]["state"][i]
            _close(st["exp_avg"].cpu().numpy(), d[f"final.{opt}.{i}.exp_avg"], what=f"{opt}[{i}].exp_avg")
            _close(st["exp_avg_sq"].cpu().numpy(), d[f"final.{opt}.{i}.exp_avg_sq"], what=f"{opt}[{i}].exp_avg_sq")
            assert float(st["step"]) == float(d[f"final.{opt}.{i}.step"])
    e = eng.export(1)
    _close(e["qt2.w0"], d["final.q_target.q2.net.0.weight"], what="target q2.w0")
    _close(e["qt1.b2"], d["final.q_target.q1.net.4.bias"], what="target q1.b2")
    assert abs(out["actor_optimizer"]["param_groups"][0]["lr"] - float(d["lr_final"])) < 1e-12
    assert out["total_it"] == int(d["total_it"]) and out["actor_lr_schedule"]["last_epoch"] == int(d["sched_last_epoch"])


def test_written_checkpoint_has_the_reference_layout(golden_dir):
    """Same updates as the reference ran before saving -> the same file, entry for entry."""
    from iqlpref_b200 import iql as biql
    d, want, hp = _fixtures(golden_dir)
    eng = biql.BatchedIQL(S, A, 1, hidden_dim=H, batch_size=B, hparams=[hp])
    eng.adopt(0, *_modules(biql, d), bind=False)
    for k in range(int(d["k_saved"])):
        eng.train(_rows(biql, d, k, eng.R).unsqueeze(0))
        _close(eng.losses[0].cpu().numpy(), d["losses"][k], what=f"losses of update {k}")
    got = eng.agent_state_dict(0, fused=None)  # the fixture was written on CPU: Adam(fused=None)
    assert list(got) == list(want)
    for part, opt in PARTS:
        assert list(got[part]) == list(want[part])  # names AND order (the optimizer state is indexed by it)
        for name in want[part]:
            assert got[part][name].dtype == want[part][name].dtype
            _close(got[part][name].cpu().numpy(), want[part][name].numpy(), what=f"{part}.{name}")
        assert sorted(got[opt]["state"]) == sorted(want[opt]["state"])
        for i, st in want[opt]["state"].items():
            assert set(got[opt]["state"][i]) == set(st)
            assert float(got[opt]["state"][i]["step"]) == float(st["step"])
            for key in ("exp_avg", "exp_avg_sq"):
                _close(got[opt]["state"][i][key].cpu().numpy(), st[key].numpy(), what=f"{opt}[{i}].{key}")
        gg, wg = got[opt]["param_groups"], want[opt]["param_groups"]
        assert len(gg) == len(wg) == 1 and set(gg[0]) == set(wg[0])
        for key, val in wg[0].items():
            if key == "lr":
                assert abs(gg[0][key] - val) < 1e-12
            else:
                assert gg[0][key] == val, key
    assert set(got["actor_lr_schedule"]) == set(want["actor_lr_schedule"])
    for key, val in want["actor_lr_schedule"].items():
        if key == "_last_lr":
            assert abs(got["actor_lr_schedule"][key][0] - val[0]) < 1e-12
        else:
            assert got["actor_lr_schedule"][key] == val, key
    assert got["total_it"] == want["total_it"]


def test_dropin_trainer_loads_the_reference_checkpoint(golden_dir):
    from iqlpref_b200 import iql as biql
    d, sd, hp = _fixtures(golden_dir)
    qf, vf, actor = (m.cuda() for m in _modules(biql, seed=5))
    mk = lambda m, lr: torch.optim.Adam(m.parameters(), lr=lr)
    tr = biql.ImplicitQLearning(1.0, actor, mk(actor, hp["actor_lr"]), qf, mk(qf, hp["qf_lr"]), vf, mk(vf, hp["vf_lr"]),
                                iql_tau=hp["iql_tau"], beta=hp["beta"], max_steps=hp["max_steps"],
                                discount=hp["discount"], tau=hp["tau"], device="cuda", batch_size=B)
    tr.load_state_dict(sd)
    assert tr.total_it == int(d["k_saved"])
    assert torch.equal(tr.qf.q2.net[2].weight.cpu(), sd["qf"]["q2.net.2.weight"])
    assert torch.equal(tr.q_target.q2.net[2].weight, tr.qf.q2.net[2].weight)
    t = {n: torch.from_numpy(d[f"table.{n}"]) for n in ("obs", "act", "rew", "next_obs", "done")}
    for k in range(int(d["k_saved"]), d["idx"].shape[0]):
        i = torch.from_numpy(d["idx"][k])
        got = tr.train([t["obs"][i].cuda(), t["act"][i].cuda(), t["rew"][i, None].cuda(), t["next_obs"][i].cuda(),
                        t["done"][i, None].cuda()])
        _close([got["value_loss"], got["q_loss"], got["actor_loss"]], d["losses"][k], what=f"update {k}")
    out = tr.state_dict()
    assert list(out["actor"]) == list(sd["actor"]) and list(out["qf"]) == list(sd["qf"])
    _close(out["actor"]["log_std"].cpu().numpy(), d["final.actor.log_std"], what="log_std")
    _close(out["actor_optimizer"]["state"][0]["exp_avg"].cpu().numpy(), d["final.actor_optimizer.0.exp_avg"], what="m(log_std)")
    assert out["total_it"] == int(d["total_it"])
    # keys written by a torch.compile'd trainer (iql.py:1523-1527) load as well
    sd2 = dict(sd, qf={"_orig_mod." + k: v for k, v in sd["qf"].items()})
    tr.load_state_dict(sd2)
    assert torch.equal(tr.qf.q1.net[0].weight.cpu(), sd["qf"]["q1.net.0.weight"])


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_per_agent_files_round_trip(tmp_path, precision):
    """save_checkpoints -> load_checkpoints into a fresh engine continues bit-identically (both precisions; on the
    tcgen05 path the bf16 operand images are rebuilt from the loaded parameters)."""
    from iqlpref_b200 import iql as biql
    N, Hh, Bb = 2, 256, 256
    hps = [dict(beta=3.0, iql_tau=0.7, max_steps=100), dict(beta=10.0, iql_tau=0.9, max_steps=40, actor_lr=1e-4)]
    rng = np.random.default_rng(3)
    batches = [torch.tensor(rng.standard_normal((N, Bb, biql.row_stride(S, A))), dtype=torch.float32).cuda() for _ in range(5)]
    for b in batches:
        b[..., 2 * S + A + 1] = (b[..., 2 * S + A + 1] > 1.5).float()

    def fresh(seed):
        eng = biql.BatchedIQL(S, A, N, hidden_dim=Hh, batch_size=Bb, hparams=hps, precision=precision)
        for a in range(N):
            torch.manual_seed(seed + a)
            eng.adopt(a, biql.TwinQ(S, A, Hh), biql.ValueFunction(S, Hh), biql.GaussianPolicy(S, A, 1.0, Hh), bind=False)
        return eng
    first = fresh(10)
    for b in batches[:3]:
        first.train(b)
    paths = [str(tmp_path / f"agent{a}" / "checkpoint_2.pt") for a in range(N)]
    for p in paths:
        os.makedirs(os.path.dirname(p))
    first.save_checkpoints(paths, total_it=3)
    second = fresh(20)
    assert second.load_checkpoints(paths) == [3, 3]
    with torch.no_grad():  # what a load does to the uninterrupted run's target (iql.py:679)
        first.target[:, :first.PQ] = first.params[:, :first.PQ]
    first._sync_shadow()
    for b in batches[3:]:
        first.train(b)
        second.train(b)
        assert torch.equal(first.losses, second.losses)
    for name in ("params", "exp_avg", "exp_avg_sq", "target", "it", "actor_lr"):
        assert torch.equal(getattr(first, name), getattr(second, name)), name
    with pytest.raises(ValueError):
        second.save_checkpoints(paths[:1])
    bad = torch.load(paths[0], weights_only=False)
    bad["vf"].pop("v.net.2.bias")
    with pytest.raises(KeyError):
        second.load_agent_state_dict(0, bad)


def test_actor_with_dropout_keeps_the_reference_key_names(golden_dir, tmp_path):
    """The reference's MLP inserts nn.Dropout after every hidden ReLU when `dropout is not None` (iql.py:436-437), so the
    actor's Linear layers are net.net.0 / 3 / 6 in its state_dict (pen / door configs, actor_dropout: 0.1).
    `tests/golden/iql_small_ckpt_dropout.pt` was written by the unmodified reference with such an actor."""
    from iqlpref_b200 import iql as biql
    sd = torch.load(os.path.join(golden_dir, "iql_small_ckpt_dropout.pt"), map_location="cpu", weights_only=False)
    assert list(sd["actor"]) == ["log_std"] + [f"net.net.{i}.{p}" for i in (0, 3, 6) for p in ("weight", "bias")]
    hp = dict(discount=0.99, tau=0.005, beta=3.0, iql_tau=0.7, vf_lr=3e-4, qf_lr=3e-4, actor_lr=3e-4, max_steps=50)
    eng = biql.BatchedIQL(S, A, 2, hidden_dim=H, batch_size=B, hparams=[hp, hp], actor_dropout=0.1)
    assert eng.load_agent_state_dict(1, sd) == 2
    e = eng.export(1)
    np.testing.assert_array_equal(e["actor.w1"], sd["actor"]["net.net.3.weight"].numpy())
    np.testing.assert_array_equal(e["actor.b2"], sd["actor"]["net.net.6.bias"].numpy())
    np.testing.assert_array_equal(eng.view(eng.exp_avg, 1, "actor.w2").cpu().numpy(),
                                  sd["actor_optimizer"]["state"][5]["exp_avg"].numpy())  # log_std, w0, b0, w1, b1, w2
    out = eng.agent_state_dict(1, fused=None)
    assert list(out["actor"]) == list(sd["actor"])  # names and order as the reference writes them
    for k, v in sd["actor"].items():
        assert torch.equal(out["actor"][k].cpu(), v)
    for i, st in sd["actor_optimizer"]["state"].items():
        assert torch.equal(out["actor_optimizer"]["state"][i]["exp_avg_sq"].cpu(), st["exp_avg_sq"])
    # save -> load round trip of the engine's own files, and an engine WITHOUT dropout modules reads the same file
    # (nn.Dropout has no parameters: only the numbering differs)
    paths = [str(tmp_path / f"a{a}.pt") for a in range(2)]
    eng.save_checkpoints(paths, total_it=2)
    assert list(torch.load(paths[1], weights_only=False)["actor"]) == list(sd["actor"])
    eng2 = biql.BatchedIQL(S, A, 2, hidden_dim=H, batch_size=B, hparams=[hp, hp], actor_dropout=0.1)
    eng2.load_checkpoints(paths)
    assert torch.equal(eng2.params[1], eng.params[1]) and torch.equal(eng2.exp_avg[1], eng.exp_avg[1])
    plain = biql.BatchedIQL(S, A, 1, hidden_dim=H, batch_size=B, hparams=[hp])
    plain.load_agent_state_dict(0, sd)
    assert torch.equal(plain.params[0], eng.params[1])
    assert list(plain.agent_state_dict(0)["actor"])[1] == "net.net.0.weight" and "net.net.2.weight" in plain.agent_state_dict(0)["actor"]
    # the drop-in trainer around a reference-shaped actor with Dropout: load, state_dict, same names
    torch.manual_seed(0)
    qf, vf = biql.TwinQ(S, A, H).cuda(), biql.ValueFunction(S, H).cuda()
    actor = biql.GaussianPolicy(S, A, 1.0, H, dropout=0.1).cuda()
    mk = lambda m, lr: torch.optim.Adam(m.parameters(), lr=lr)
    tr = biql.ImplicitQLearning(1.0, actor, mk(actor, 3e-4), qf, mk(qf, 3e-4), vf, mk(vf, 3e-4), max_steps=50,
                                device="cuda", batch_size=B)
    assert list(tr.state_dict()["actor"]) == list(sd["actor"])  # the advisor's case: the trainer's own state_dict loads
    tr.load_state_dict(tr.state_dict())
    tr.load_state_dict(sd)
    assert tr.total_it == 2
    assert torch.equal(tr.actor.net.net[3].weight.cpu(), sd["actor"]["net.net.3.weight"])
    got = tr.state_dict()
    assert list(got["actor"]) == list(sd["actor"])
    assert torch.equal(got["actor_optimizer"]["state"][3]["exp_avg"].cpu(), sd["actor_optimizer"]["state"][3]["exp_avg"])


def test_actor_lr_schedule_attribute(golden_dir):
    """`trainer.actor_lr_schedule` (iql.py:571): get_last_lr / state_dict / load_state_dict over the in-kernel schedule."""
    from iqlpref_b200 import iql as biql
    d, sd, hp = _fixtures(golden_dir)
    qf, vf, actor = (m.cuda() for m in _modules(biql, d))
    mk = lambda m, lr: torch.optim.Adam(m.parameters(), lr=lr)
    tr = biql.ImplicitQLearning(1.0, actor, mk(actor, hp["actor_lr"]), qf, mk(qf, hp["qf_lr"]), vf, mk(vf, hp["vf_lr"]),
                                iql_tau=hp["iql_tau"], beta=hp["beta"], max_steps=hp["max_steps"],
                                discount=hp["discount"], tau=hp["tau"], device="cuda", batch_size=B)
    sch = tr.actor_lr_schedule
    assert sch.get_last_lr() == [hp["actor_lr"]] and sch.last_epoch == 0 and sch.T_max == hp["max_steps"]
    t = {n: torch.from_numpy(d[f"table.{n}"]) for n in ("obs", "act", "rew", "next_obs", "done")}
    for k in range(int(d["k_saved"])):
        i = torch.from_numpy(d["idx"][k])
        tr.train([t["obs"][i].cuda(), t["act"][i].cuda(), t["rew"][i, None].cuda(), t["next_obs"][i].cuda(),
                  t["done"][i, None].cuda()])
    assert sch.last_epoch == int(d["k_saved"])
    assert abs(sch.get_last_lr()[0] - float(d["lr_saved"])) < 1e-15  # the reference's scheduler after the same steps
    want = sd["actor_lr_schedule"]
    got = sch.state_dict()
    assert set(got) == set(want)
    for key, val in want.items():
        if key == "_last_lr":
            assert abs(got[key][0] - val[0]) < 1e-15
        else:
            assert got[key] == val, key
    sch.load_state_dict(dict(want, _last_lr=[1.25e-4]))
    assert sch.get_last_lr() == [1.25e-4] and tr.actor_optimizer.param_groups[0]["lr"] == 1.25e-4
    with pytest.raises(RuntimeError):
        sch.step()
