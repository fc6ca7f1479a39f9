"""clock64 timeline of CTA 0 of the tensor-core relabel kernel (antmaze q_mlp shape): python scripts/relabel_timeline.py"""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from iqlpref_b200 import _lib  # noqa: E402
from iqlpref_b200.relabel import ensemble_predict, pack_members  # noqa: E402

S = int(os.environ.get("TL_MEMBERS", "1"))
rng = np.random.default_rng(0)
members = []
for _ in range(S):
    m, n_in = [], 37
    for _l in range(3):
        m += [rng.standard_normal((n_in, 128), dtype=np.float32), rng.standard_normal(128, dtype=np.float32)]
        n_in = 128
    m += [rng.standard_normal((128, 1), dtype=np.float32), rng.standard_normal(1, dtype=np.float32)]
    members.append(m)
packed = pack_members(members)
table = torch.randn((1_000_000, 68), device="cuda")
x = table[:, :37]
for _ in range(3):
    ensemble_predict(packed, x, "mean", precision="split_bf16")
buf = torch.zeros(5 * 512, dtype=torch.int64, device="cuda")
lib = _lib.load()
lib.iqlb_relabel_debug_timeline(C.c_void_p(buf.data_ptr()))
ensemble_predict(packed, x, "mean", precision="split_bf16")
torch.cuda.synchronize()
lib.iqlb_relabel_debug_timeline(None)
t = buf.cpu().numpy().reshape(5, 512)
t0 = t[t > 0].min()
ev = []
for i in range(6):
    for l in range(3):
        for p in range(2):
            k = ((i * 3 + l) * 2 + p) * 2
            if t[0, k]:
                ev.append((t[0, k] - t0, f"MMA  i{i} L{l} p{p} issue  (+{t[0, k + 1] - t[0, k]} clk to commit)"))
    for role in range(1, 5):
        p, half = (role - 1) % 2, (role - 1) // 2
        for st in range(4):
            k = (i * 4 + st) * 2
            if t[role, k]:
                name = "X->TMEM" if st == 0 else f"EPI L{st - 1}"
                ev.append((t[role, k] - t0, f"  p{p} h{half} i{i} {name} start (+{t[role, k + 1] - t[role, k]} clk)"))
for c, s in sorted(ev):
    print(f"{c:8d}  {s}")
