"""GPU parity: replay sampling (index stream bit-exact with torch.randint on CUDA; rows bit-exact)."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle.philox_oracle import cuda_randint_indices

pytestmark = pytest.mark.gpu


def _lib():
    from iqlpref_b200 import _lib
    return _lib


@pytest.mark.parametrize("seed,offset,n,batch", [(0, 0, 1_000_000, 256), (7, 4, 999_983, 256), (123456789012, 40, 5000, 64),
                                                 (3, 0, 1, 5), (5, 8, 17, 1), (1, 0, (1 << 28) + 12345, 256),
                                                 (2, 12, 1_000_000, 1000)])
def test_philox_randint_matches_oracle_and_torch(seed, offset, n, batch):
    L = _lib()
    lib = L.load()
    out = torch.empty(batch, dtype=torch.int64, device="cuda")
    L.check(lib.iqlb_philox_randint(seed, offset, n, batch, L.ptr(out), L.stream_ptr()), "randint")
    got = out.cpu().numpy()
    np.testing.assert_array_equal(got, cuda_randint_indices(seed, offset, n, batch))
    # the real thing: torch's CUDA generator at (seed, offset)
    gen = torch.cuda.default_generators[torch.cuda.current_device()]
    torch.manual_seed(seed)
    gen.set_offset(offset)
    want = torch.randint(0, n, (batch,), device="cuda").cpu().numpy()
    np.testing.assert_array_equal(got, want)


def _fill(buf, rng, n, S, A):
    data = {"observations": rng.standard_normal((n, S)).astype(np.float32),
            "actions": rng.uniform(-1, 1, (n, A)).astype(np.float32),
            "rewards": rng.standard_normal(n).astype(np.float32),
            "next_observations": rng.standard_normal((n, S)).astype(np.float32),
            "terminals": (rng.uniform(size=n) < 0.1)}
    buf.load_d4rl_dataset(data)
    return data


@pytest.mark.parametrize("S,A", [(17, 6), (29, 8), (45, 24)])
def test_replay_buffer_sample_is_dropin(S, A):
    from iqlpref_b200.iql import ReplayBuffer
    rng = np.random.default_rng(0)
    n = 10_000
    buf = ReplayBuffer(S, A, 20_000, "cuda")
    data = _fill(buf, rng, n, S, A)
    ref = {k: torch.tensor(np.asarray(v, dtype=np.float32), device="cuda") for k, v in data.items()}
    for seed in (0, 11):
        torch.manual_seed(seed)
        i1 = torch.randint(0, n, (256,), device="cuda")
        i2 = torch.randint(0, n, (256,), device="cuda")
        torch.manual_seed(seed)
        b1 = buf.sample(256)
        b2 = buf.sample(256)
        for idx, b in ((i1, b1), (i2, b2)):
            assert [tuple(t.shape) for t in b] == [(256, S), (256, A), (256, 1), (256, S), (256, 1)]
            assert torch.equal(b[0], ref["observations"][idx])
            assert torch.equal(b[1], ref["actions"][idx])
            assert torch.equal(b[2], ref["rewards"][idx][:, None])
            assert torch.equal(b[3], ref["next_observations"][idx])
            assert torch.equal(b[4], ref["terminals"][idx][:, None])
        # generator state advanced exactly like two torch.randint calls
        torch.manual_seed(seed)
        torch.randint(0, n, (256,), device="cuda"); torch.randint(0, n, (256,), device="cuda")
        want_next = torch.randint(0, n, (256,), device="cuda")
        torch.manual_seed(seed)
        buf.sample(256); buf.sample(256)
        got_next = torch.randint(0, n, (256,), device="cuda")
        assert torch.equal(want_next, got_next)


def test_replay_errors():
    from iqlpref_b200.iql import ReplayBuffer
    buf = ReplayBuffer(3, 2, 10, "cuda")
    rng = np.random.default_rng(0)
    with pytest.raises(RuntimeError):
        buf.sample(4)  # empty
    with pytest.raises(ValueError):
        _fill(ReplayBuffer(3, 2, 5, "cuda"), rng, 6, 3, 2)
    _fill(buf, rng, 10, 3, 2)
    with pytest.raises(ValueError):
        _fill(buf, rng, 2, 3, 2)
    with pytest.raises(NotImplementedError):
        buf.add_transition()
    with pytest.raises(RuntimeError):
        ReplayBuffer(3, 2, 10, "cpu")


def test_multi_agent_sample_and_reward_variants():
    L = _lib()
    lib = L.load()
    from iqlpref_b200.iql import ReplayBuffer
    S, A, n, N, B = 5, 3, 777, 4, 96
    rng = np.random.default_rng(1)
    buf = ReplayBuffer(S, A, n, "cuda")
    _fill(buf, rng, n, S, A)
    variants = torch.tensor(rng.standard_normal((3, n)).astype(np.float32), device="cuda")
    which = torch.tensor([2, 0, 1, 2], dtype=torch.int32, device="cuda")
    state = torch.tensor([[10, 0], [11, 4], [12, 8], [10, 0]], dtype=torch.int64, device="cuda")
    idx = torch.empty((N, B), dtype=torch.int64, device="cuda")
    out = torch.empty((N, B, buf._stride), dtype=torch.float32, device="cuda")
    rc = 2 * S + A
    L.check(lib.iqlb_replay_sample(L.ptr(buf._table), n, buf._stride, rc, L.ptr(variants), L.ptr(which), L.ptr(state),
                                   N, B, L.ptr(idx), L.ptr(out), L.stream_ptr()), "sample")
    idx_h = idx.cpu().numpy()
    for a in range(N):
        s, o = state[a].tolist()
        np.testing.assert_array_equal(idx_h[a], cuda_randint_indices(s, o, n, B))
        want = buf._table[idx[a]].clone()
        want[:, rc] = variants[which[a].item(), idx[a]]
        assert torch.equal(out[a], want)
    # explicit-index gather gives the same rows
    out2 = torch.empty_like(out)
    L.check(lib.iqlb_replay_gather(L.ptr(buf._table), n, buf._stride, rc, L.ptr(variants), L.ptr(which), L.ptr(idx), N, B,
                                   L.ptr(out2), L.stream_ptr()), "gather")
    assert torch.equal(out, out2)


def test_polyak_matches_lerp():
    L = _lib()
    lib = L.load()
    for n in (1, 3, 4, 1023, 151554):
        t = torch.randn(n, device="cuda")
        s = torch.randn(n, device="cuda")
        want = t.clone().lerp_(s, 0.005)
        L.check(lib.iqlb_polyak(L.ptr(t), L.ptr(s), n, 0.005, L.stream_ptr()), "polyak")
        torch.testing.assert_close(t, want, rtol=1e-6, atol=1e-7)
