"""tcgen05 issue-rate probe (one CTA): clocks per 128 x n x 16 bf16 MMA, A from shared memory vs tensor memory."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from iqlpref_b200 import _lib  # noqa: E402

lib = _lib.load()
out = torch.zeros(2, dtype=torch.int64, device="cuda")
print("| A operand | N | MMAs | clocks/MMA (issue) | clocks/MMA (to completion) | dense bf16 flop/clk/SM |")
print("|---|---:|---:|---:|---:|---:|")
for mode in (0, 1, 2, 3, 4, 5, 6, 7):
    for n in ((64, 128, 256) if not mode & 2 else (64, 128)):
        for count in (64, 512):
            for _ in range(2):
                _lib.check(lib.iqlb_selftest_mma_rate(mode, n, count, C.c_void_p(out.data_ptr()), None), "mma_rate")
                torch.cuda.synchronize()
            t_issue, t_all = out.tolist()
            iss = 2 if mode & 2 else 1
            print(f"| {'TMEM' if mode & 1 else 'SMEM'}{' x2 issuers' if iss == 2 else ''}{' (elect.sync)' if mode & 4 else ''} | {n} | {count * iss} | "
                  f"{t_issue / count:.1f} | {t_all / (count * iss):.1f} | {128 * n * 16 * 2 * count * iss / t_all:.0f} |")
