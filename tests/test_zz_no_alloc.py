"""`iqlb_reward_range_static` (include/iqlb.h): the allocation-free variant of `iqlb_reward_range`, its scan scratch in a
static device array of the library.  No counterpart in the reference, so the
checker is the existing entry point: same inputs through the C ABI, results bit for bit (the bar of
tests/test_gpu_dataset_prep.py: min / max return and episode lengths are exact), on two calls in a row with different
inputs (the second call sees the first call's scratch contents).  Collected last on
purpose."""
import ctypes as C

import numpy as np
import pytest
import torch

from iqlpref_b200 import _lib

pytestmark = pytest.mark.gpu

MAX_ROWS = 16777216  # IQLB_REWARD_RANGE_STATIC_MAX_ROWS


def _dataset(seed, n, p_term):
    rng = np.random.default_rng(seed)
    r = torch.from_numpy(rng.standard_normal(n).astype(np.float32)).cuda()
    t = torch.from_numpy((rng.uniform(size=n) < p_term).astype(np.uint8)).cuda()
    return r, t


def _call(name, r, t, max_steps, want_lens=True):
    lib = _lib.load()
    dev = r.device
    mm = torch.full((2,), float("nan"), dtype=torch.float64, device=dev)
    lens = torch.full((r.numel(),), -1.0, dtype=torch.float32, device=dev) if want_lens else None
    with _lib.device_ctx(dev):
        rc = getattr(lib, name)(_lib.ptr(r), _lib.ptr(t), r.numel(), max_steps, _lib.ptr(mm), _lib.ptr(lens),
                                _lib.stream_ptr(dev))
    _lib.check(rc, name)
    torch.cuda.synchronize()
    return mm.cpu().numpy(), (lens.cpu().numpy() if want_lens else None)


def test_static_scratch_variant_matches_the_allocating_entry_point_twice_in_a_row():
    # first a large ragged dataset (fills ~600 scratch entries), then a small one with other episode boundaries
    for seed, n, p_term, max_steps in ((0, 600_011, 0.002, 700), (1, 4_099, 0.01, 57)):
        r, t = _dataset(seed, n, p_term)
        want_mm, want_lens = _call("iqlb_reward_range", r, t, max_steps)
        got_mm, got_lens = _call("iqlb_reward_range_static", r, t, max_steps)
        assert np.isfinite(want_mm).all()
        assert np.array_equal(got_mm, want_mm), (got_mm, want_mm)
        assert np.array_equal(got_lens, want_lens)
    # trj_lens is nullable, as in the allocating entry point
    got_mm, _ = _call("iqlb_reward_range_static", r, t, max_steps, want_lens=False)
    assert np.array_equal(got_mm, want_mm)


def test_static_scratch_variant_rejects_sizes_beyond_its_maximum():
    lib = _lib.load()
    r, t = _dataset(2, 16, 0.1)
    mm = torch.zeros(2, dtype=torch.float64, device="cuda")
    rc = lib.iqlb_reward_range_static(_lib.ptr(r), _lib.ptr(t), MAX_ROWS + 1, 10, _lib.ptr(mm), None, None)
    assert rc != 0  # refused before any launch: the 16-row buffers are never read
    with pytest.raises(ValueError, match="at most"):  # IQLB_ERR_INVALID surfaces as ValueError (_lib.check)
        _lib.check(rc, "iqlb_reward_range_static")
    assert lib.iqlb_reward_range_static(None, _lib.ptr(t), 16, 10, _lib.ptr(mm), None, None) != 0
