#!/usr/bin/env python
"""Microbenchmark of the generic GEMM (tensor-core split-bf16 vs fp32 FFMA) through iqlb_selftest_gemm."""
import ctypes as C, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
from iqlpref_b200 import _lib

lib = _lib.load()
dev = torch.device("cuda", 0)
for (m, n, k) in [(192000, 512, 256), (192000, 512, 64), (192000, 512, 1024), (192000, 128, 256), (16384, 1024, 256), (51200, 128, 128)]:
    a = torch.randn((m, k), device=dev); w = torch.randn((n, k), device=dev) / k ** 0.5
    b = torch.zeros(n, device=dev); c = torch.empty((m, n), device=dev)
    res = {"m": m, "n": n, "k": k}
    for tc in (1, 0):
        fn = lambda: _lib.check(lib.iqlb_selftest_gemm(0, _lib.ptr(a), k, _lib.ptr(w), k, _lib.ptr(b), None, _lib.ptr(c), n, m, n, k,
                                                       4, C.c_float(1.0), tc, _lib.stream_ptr(dev)), "gemm")
        fn(); fn()
        ms = min(bench.time_gpu(fn, 5, dev) for _ in range(2))
        res["tc" if tc else "ffma"] = {"ms": round(ms, 4), "tflops_fp32_equiv": round(2.0 * m * n * k / (ms * 1e-3) / 1e12, 1)}
    print(json.dumps(res))
