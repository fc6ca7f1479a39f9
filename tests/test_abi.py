"""CPU: the C-ABI library loads and exports every symbol include/iqlb.h declares; host-only entry
points (no GPU work) behave."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from oracle import relabel_oracle as ro

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from iqlpref_b200 import _lib
    if not os.path.isfile(_lib.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    return _lib.load()


def test_every_declared_symbol_is_exported(lib):
    from iqlpref_b200 import _lib
    header = open(os.path.join(ROOT, "include", "iqlb.h")).read()
    declared = set(re.findall(r"\b(iqlb_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations found"
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in iqlb.h but not exported by libiqlb.so"
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)


def test_param_count_matches_reference_shapes(lib):
    # 286 991 / 300 819 / 329 523 params per agent (Gaussian), Q = 144 386 / 151 554 / 167 938;
    # the packed layout pads each of the four network segments to a multiple of 8 floats
    al8 = lambda n: (n + 7) // 8 * 8
    for (s, a), (p, pq) in {(17, 6): (286991, 144386), (29, 8): (300819, 151554), (45, 24): (329523, 167938)}.items():
        qnet, vnet, anet = pq // 2, 256 * s + 256 + 65536 + 256 + 256 + 1, p - a - pq - (256 * s + 256 + 65536 + 256 + 256 + 1)
        assert qnet * 2 == pq and anet == 256 * s + 256 + 65536 + 256 + 256 * a + a
        pa, pb = C.c_int64(), C.c_int64()
        assert lib.iqlb_param_count(s, a, 256, 0, C.byref(pa), C.byref(pb)) == 0
        assert (pa.value, pb.value) == (2 * al8(qnet) + al8(vnet) + al8(anet) + a, 2 * al8(qnet))
        assert lib.iqlb_param_count(s, a, 256, 1, C.byref(pa), C.byref(pb)) == 0
        assert pa.value == 2 * al8(qnet) + al8(vnet) + al8(anet)
    assert lib.iqlb_param_count(0, 1, 256, 0, None, None) == -1
    assert b"dims" in lib.iqlb_last_error()


@pytest.mark.parametrize("with_timeouts", [True, False])
@pytest.mark.parametrize("toe", [False, True])
def test_keep_mask_scan_matches_oracle(lib, with_timeouts, toe):
    rng = np.random.default_rng(0)
    n = 5000
    term = (rng.uniform(size=n) < 0.01).astype(np.uint8)
    tout = np.zeros(n, np.uint8)
    tout[99::100] = 1
    keep = np.zeros(n - 1, np.uint8)
    steps = np.zeros(n - 1, np.int32)
    rc = lib.iqlb_keep_mask_scan(term.ctypes.data_as(C.c_void_p), tout.ctypes.data_as(C.c_void_p) if with_timeouts else None,
                                 n, 37, int(toe), keep.ctypes.data_as(C.c_void_p), steps.ctypes.data_as(C.c_void_p))
    assert rc == 0
    k2, s2 = ro.keep_mask_scan(term, tout if with_timeouts else None, 37, toe)
    np.testing.assert_array_equal(keep.astype(bool), k2)
    np.testing.assert_array_equal(steps, s2)


def test_host_mirror_surface():
    from iqlpref_b200 import iql as biql
    cfg = biql.TrainConfig(seed=3, reward_model_root="/x/model")
    assert cfg.reward_model_path == "/x/model_3" and cfg.batch_size == 256 and cfg.mr_alpha == 0.95
    assert biql.row_stride(29, 8) == 68 and biql.row_stride(17, 6) == 44 and biql.row_stride(45, 24) == 116
    with pytest.raises(ValueError):
        biql.MLP([3])
    with pytest.raises(ValueError):
        biql.MLP([3, 4, 2], squeeze_output=True)
    sd = biql.TwinQ(5, 2, 8).state_dict()
    assert list(sd)[:2] == ["q1.net.0.weight", "q1.net.0.bias"] and "q2.net.4.bias" in sd
    assert "log_std" in biql.GaussianPolicy(5, 2, 1.0, 8).state_dict()
    from iqlpref_b200.relabel import cvar_n_tail, pt_windows_from_steps
    assert cvar_n_tail(0.9, 40) == 3 and cvar_n_tail(0.9, 10) == 1 and cvar_n_tail(0.0, 7) == 7
    st, sl = pt_windows_from_steps(np.array([0, 5, 9, 10, 57]), 10)
    assert st.tolist() == [0, 0, 0, 1, 48] and sl.tolist() == [1, 6, 10, 10, 10]


def test_mr_trainer_host_surface():
    """MRTrainer mirror (training.py:11-129): argument checks that need no GPU; there is no CPU path."""
    import torch
    from iqlpref_b200.mr_training import MRTrainer, fit_mr
    from iqlpref_b200.reward_models.q_mlp import MLP
    net = MLP(6, 1, [8, 8], "relu")
    opt = torch.optim.Adam(net.parameters(), lr=3e-4)
    with pytest.raises(RuntimeError, match="no CPU path"):
        MRTrainer(net, opt, num_datapoints=10, device="cpu")
    with pytest.raises(ValueError, match="bt_pool_logit"):
        MRTrainer(net, opt, num_datapoints=10, device="cuda", bt_pool="median")
    with pytest.raises(NotImplementedError):
        MRTrainer(net, torch.optim.SGD(net.parameters(), lr=0.1), num_datapoints=10, device="cuda")
    with pytest.raises(NotImplementedError):
        MRTrainer(net, torch.optim.Adam(net.parameters(), lr=1e-3, weight_decay=0.1), num_datapoints=10, device="cuda")
    with pytest.raises(ValueError, match="criteria_key"):
        fit_mr(None, [], [], epochs=1, criteria_key="f1")
    assert list(net.state_dict())[:4] == ["layers.0.W", "layers.0.b", "layers.linear_1.W", "layers.linear_1.b"]


def test_mr_entry_points_reject_bad_arguments_without_touching_the_gpu(lib):
    h = C.c_void_p()
    widths = (C.c_int32 * 2)(8, 8)
    assert lib.iqlb_mr_create(0, widths, 2, 0, 1, 64, C.byref(h)) == -1 and b"network shape" in lib.iqlb_last_error()
    assert lib.iqlb_mr_create(6, widths, 2, 99, 1, 64, C.byref(h)) == -1 and b"activation" in lib.iqlb_last_error()
    assert lib.iqlb_mr_create(6, widths, 2, 0, 1, 63, C.byref(h)) == -1 and b"max_rows" in lib.iqlb_last_error()
    assert lib.iqlb_mr_create(6, widths, 2, 0, 1, 64, None) == -1
    assert lib.iqlb_mr_param_count(None, None) == -1
    assert lib.iqlb_mr_train_step(None, *([None] * 10), 1, 1, 1, 1, 0.0, 0.9, 0.999, 1e-8, 1, None, None, None) == -1
    assert lib.iqlb_mr_eval(None, *([None] * 8), 1, 1, 1, 1, None, None, None) == -1
    assert lib.iqlb_mr_destroy(None) == 0
