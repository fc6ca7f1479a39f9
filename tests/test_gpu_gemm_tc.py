"""The generic split-bf16 tensor-core GEMM (csrc/gemm_tc.cu) against numpy float64, through the library's dispatcher.
Tolerance: every product carries ~2^-17 relative error (three bf16 MMAs, fp32 accumulation in TMEM), so an output is
within 2e-5 of the row's |a| . |w| scale; the fp32 FFMA kernel on the same inputs is the second reference."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ACT = {"relu": 0, "tanh": 1, "linear": 4}


def _gemm(kn, a, w, bias, residual, n, k, act, scale, allow_tc, ldc=None):
    from iqlpref_b200 import _lib
    lib = _lib.load()
    m = a.shape[0]
    ldc = ldc or n
    c = torch.zeros((m, ldc), dtype=torch.float32, device="cuda")
    _lib.check(lib.iqlb_selftest_gemm(int(kn), _lib.ptr(a), a.stride(0), _lib.ptr(w), w.stride(0), _lib.ptr(bias),
                                      _lib.ptr(residual), _lib.ptr(c), ldc, m, n, k, ACT[act], C.c_float(scale),
                                      int(allow_tc), _lib.stream_ptr(torch.device("cuda"))), "iqlb_selftest_gemm")
    return c[:, :n]


@pytest.mark.parametrize("kn", [0, 1])
@pytest.mark.parametrize("m,n,k,act", [(128, 128, 64, "linear"), (1000, 512, 256, "relu"), (513, 200, 37, "tanh"),
                                        (9600, 768, 256, "linear"), (257, 64, 1024, "relu"), (130, 129, 17, "linear")])
def test_tensor_core_gemm_matches_numpy(kn, m, n, k, act):
    rng = np.random.default_rng(m + n + k + kn)
    a_h = rng.standard_normal((m, k)).astype(np.float32)
    w_h = (rng.standard_normal((k, n) if kn else (n, k)) / np.sqrt(k)).astype(np.float32)
    b_h = rng.standard_normal(n).astype(np.float32)
    r_h = rng.standard_normal((m, n)).astype(np.float32)
    scale = 0.7 if kn else 1.0
    a, w, b, r = (torch.from_numpy(x).cuda() for x in (a_h, w_h, b_h, r_h))
    acc = a_h.astype(np.float64) @ (w_h.astype(np.float64) if kn else w_h.astype(np.float64).T)
    z = scale * acc + b_h
    want = {"linear": z, "relu": np.maximum(z, 0), "tanh": np.tanh(z)}[act]
    if not kn:
        want = want + r_h
    bound = 2e-5 * (np.abs(a_h).astype(np.float64) @ np.abs(w_h.astype(np.float64) if kn else w_h.T.astype(np.float64))) + 2e-6
    got_tc = _gemm(kn, a, w, b, None if kn else r, n, k, act, scale, 1).cpu().numpy()
    got_ff = _gemm(kn, a, w, b, None if kn else r, n, k, act, scale, 0).cpu().numpy()
    assert (np.abs(got_ff - want) <= bound).all(), "FFMA reference path"
    err = np.abs(got_tc - want)
    assert (err <= bound).all(), f"max err {err.max():.3e} at {np.unravel_index(err.argmax(), err.shape)}"


def test_tensor_core_gemm_strided_and_unaligned_operands():
    """Row-strided A (a column slice of a wider matrix, rows not 16-byte aligned), padded output rows."""
    rng = np.random.default_rng(5)
    m, n, k = 300, 256, 100
    wide = torch.from_numpy(rng.standard_normal((m, k + 7)).astype(np.float32)).cuda()
    a = wide[:, 3:3 + k]
    w = torch.from_numpy((rng.standard_normal((n, k + 1)) / 10).astype(np.float32)).cuda()[:, :k]
    b = torch.zeros(n, device="cuda")
    want = a.double().cpu().numpy() @ w.double().cpu().numpy().T
    got = _gemm(0, a, w, b, None, n, k, "linear", 1.0, 1, ldc=n + 3).cpu().numpy()
    np.testing.assert_allclose(got, want, rtol=0, atol=3e-5 * np.abs(want).max() + 1e-5)
