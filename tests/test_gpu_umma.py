"""GPU: the tcgen05 building blocks (shared-memory descriptors for K-major and MN-major operand
images, instruction descriptor, TMEM read-back) against numpy, for every operand arrangement of
the IQL step: forward (A K-major, B K-major), data gradient (A K-major, B MN-major) and weight
gradient (A MN-major, B MN-major).  bf16 operands, fp32 accumulate: tolerance 1e-3 of |D|max on
bf16-rounded inputs (only the fp32 summation order differs)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def to_image(mat: torch.Tensor) -> torch.Tensor:
    """[R, C] float -> bf16 operand image (uint8 bytes) in the layout of csrc/umma.cuh."""
    r, c = mat.shape
    cb = (c + 63) // 64
    m = torch.zeros((r, cb * 64), dtype=torch.bfloat16)
    m[:, :c] = mat.to(torch.bfloat16)
    m = m.view(r, cb, 8, 8)                       # [row, block, chunk, elem]
    rows = torch.arange(r).view(r, 1, 1) % 8
    chunk = torch.arange(8).view(1, 1, 8)
    src = (chunk ^ rows).expand(r, cb, 8)         # physical chunk p holds logical chunk p ^ (row % 8)
    img = torch.gather(m, 2, src.unsqueeze(-1).expand(r, cb, 8, 8))
    return img.permute(1, 0, 2, 3).contiguous().view(torch.uint8).reshape(-1)


def _run(a_img, b_img, a_mn, a_rows, a_off, b_mn, b_rows, b_off, n, ksteps):
    from iqlpref_b200 import _lib
    lib = _lib.load()
    a, b = a_img.cuda(), b_img.cuda()
    out = torch.full((128, n), float("nan"), dtype=torch.float32, device="cuda")
    _lib.check(lib.iqlb_selftest_umma(_lib.ptr(a), a.numel(), _lib.ptr(b), b.numel(), a_mn, a_rows, a_off, b_mn, b_rows,
                                      b_off, n, ksteps, _lib.ptr(out), _lib.stream_ptr()), "iqlb_selftest_umma")
    torch.cuda.synchronize()
    return out.cpu()


def _bf(x):
    return x.to(torch.bfloat16).to(torch.float32)


def _check(got, want):
    scale = float(want.abs().max())
    err = float((got - want).abs().max())
    assert torch.isfinite(got).all()
    assert err <= 1e-3 * scale, f"max err {err:.3e} vs scale {scale:.3e}"


@pytest.mark.parametrize("k,n", [(256, 256), (48, 256), (80, 256), (256, 32), (16, 16)])
def test_forward_arrangement(k, n):
    g = torch.Generator().manual_seed(k * 1000 + n)
    x, w = torch.randn(128, k, generator=g), torch.randn(n, k, generator=g)
    got = _run(to_image(x), to_image(w), 0, 128, 0, 0, n, 0, n, (k + 15) // 16)
    _check(got, _bf(x) @ _bf(w).T)


def test_forward_row_offset_into_a_larger_image():
    g = torch.Generator().manual_seed(7)
    x, w = torch.randn(256, 128, generator=g), torch.randn(256, 128, generator=g)  # 64 KB + 64 KB of smem
    got = _run(to_image(x), to_image(w), 0, 256, 128, 0, 256, 0, 256, 8)
    _check(got, _bf(x[128:]) @ _bf(w).T)


@pytest.mark.parametrize("n", [256, 64])
def test_data_gradient_arrangement(n):
    g = torch.Generator().manual_seed(11 + n)
    dy, w = torch.randn(128, 256, generator=g), torch.randn(256, n, generator=g)  # W[out=K][in=N]
    got = _run(to_image(dy), to_image(w), 0, 128, 0, 1, 256, 0, n, 16)
    _check(got, _bf(dy) @ _bf(w))


@pytest.mark.parametrize("n_in,col0", [(128, 0), (128, 128), (48, 128), (80, 0)])
def test_weight_gradient_arrangement(n_in, col0):
    g = torch.Generator().manual_seed(13 + n_in + col0)
    dy, x = torch.randn(256, 256, generator=g), torch.randn(256, n_in, generator=g)
    n = (n_in + 15) // 16 * 16
    got = _run(to_image(dy), to_image(x), 1, 256, col0, 1, 256, 0, n, 16)
    want = _bf(dy[:, col0:col0 + 128]).T @ _bf(x)
    _check(got[:, :n_in], want)
