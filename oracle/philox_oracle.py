"""CPU oracle for the replay-buffer index stream (TEST INFRASTRUCTURE, not product).

Restates what `torch.randint(0, n, (B,), device="cuda")` computes on the reference's
sampling path (`/root/reference/algorithms/offline/iql.py:211-214`).  The arithmetic
lives in third-party code that is not under /root/reference:

* PyTorch 2.11 `ATen/native/cuda/DistributionTemplates.h:50-63` (execution policy),
  `:66-90` (grid-stride kernel, `curand_init(seed, idx, offset)`), `:319-345`
  (32-bit path `rand.x % range` for range < 2**28, else `((x<<32)|y) % range`).
* cuRAND `curand_kernel.h` (`curand_init` for Philox4_32_10: ctr=0, key=seed,
  `skipahead_sequence(subsequence)` adds to ctr.zw, `skipahead(offset)` adds offset/4
  to ctr.xy) and `curand_philox4x32_x.h` (`curand_Philox4x32_10`, 10 rounds,
  multipliers 0xD2511F53 / 0xCD9E8D57, Weyl keys 0x9E3779B9 / 0xBB67AE85).

Pinned by (a) the Random123 known-answer vectors for Philox4x32-10 (tests/test_philox.py)
and (b) on the GPU box, bit-exact comparison with `torch.randint(..., device="cuda")`.
"""
import numpy as np

M0 = np.uint64(0xD2511F53)
M1 = np.uint64(0xCD9E8D57)
W0 = 0x9E3779B9
W1 = 0xBB67AE85
MASK32 = np.uint64(0xFFFFFFFF)


def philox4x32_10(ctr, key):
    """ctr: (..., 4) uint32-valued, key: (..., 2) uint32-valued -> (..., 4) uint32."""
    c = [np.asarray(ctr[..., i], dtype=np.uint64) for i in range(4)]
    k0 = np.asarray(key[..., 0], dtype=np.uint64)
    k1 = np.asarray(key[..., 1], dtype=np.uint64)
    for r in range(10):
        p0 = M0 * c[0]
        p1 = M1 * c[2]
        hi0, lo0 = p0 >> np.uint64(32), p0 & MASK32
        hi1, lo1 = p1 >> np.uint64(32), p1 & MASK32
        c = [hi1 ^ c[1] ^ k0, lo1, hi0 ^ c[3] ^ k1, lo0]
        k0 = (k0 + np.uint64(W0)) & MASK32
        k1 = (k1 + np.uint64(W1)) & MASK32
    return np.stack(c, axis=-1).astype(np.uint32)


def cuda_randint_indices(seed: int, offset: int, n: int, batch: int) -> np.ndarray:
    """Indices `torch.randint(0, n, (batch,), device='cuda')` yields for a generator at
    (seed, philox offset).  Valid while batch <= 256 * grid (every thread emits only its
    first draw), i.e. for every batch the IQL path uses."""
    if offset % 4 != 0:
        raise ValueError("torch philox offsets are multiples of 4")
    t = np.arange(batch, dtype=np.uint64)
    blk = np.uint64(offset // 4)
    ctr = np.stack(
        [
            np.full(batch, blk & MASK32, dtype=np.uint64),
            np.full(batch, blk >> np.uint64(32), dtype=np.uint64),
            t & MASK32,
            t >> np.uint64(32),
        ],
        axis=-1,
    )
    key = np.empty((batch, 2), dtype=np.uint64)
    key[:, 0] = np.uint64(seed & 0xFFFFFFFF)
    key[:, 1] = np.uint64((seed >> 32) & 0xFFFFFFFF)
    out = philox4x32_10(ctr, key).astype(np.uint64)
    if n >= (1 << 28):
        r = (out[:, 0] << np.uint64(32)) | out[:, 1]
    else:
        r = out[:, 0]
    return (r % np.uint64(n)).astype(np.int64)


def offset_increment(batch: int, n: int, n_sms: int = 148, max_threads_per_sm: int = 2048) -> int:
    """Generator offset advance per randint call (DistributionTemplates.h:50-63)."""
    block = 256
    grid = min(n_sms * (max_threads_per_sm // block), (batch + block - 1) // block)
    unroll = 2 if n >= (1 << 28) else 4
    return ((batch - 1) // (block * grid * unroll) + 1) * 4
