#!/usr/bin/env python
"""PT relabel timing (events around pt_predict, weights cached): python scripts/bench_pt.py [n_windows] [layers]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import bench
from iqlpref_b200.relabel import pt_predict
from iqlpref_b200.reward_models.pref_transformer import PT

n_win = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
L = int(sys.argv[2]) if len(sys.argv) > 2 else 1
dev = torch.device("cuda", 0)
rows, q = 200_000, 100
torch.manual_seed(0)
pt = PT(state_dim=29, action_dim=8, max_episode_steps=1000, embd_dim=256, pref_attn_embd_dim=256, num_heads=4,
        intermediate_dim=1024, num_layers=L, max_pos=2048, eps=0.1).to(dev).eval()
obs = torch.randn((rows, 29), device=dev)
act = torch.rand((rows, 8), device=dev) * 2 - 1
start = (np.arange(n_win, dtype=np.int64) * 7 % (rows - q)).astype(np.int32)
seq = np.full(n_win, q, dtype=np.int32)
seq[::10] = 1 + np.arange(len(seq[::10])) % q
fn = lambda: pt_predict(pt, obs, act, start, seq, q, device=dev, dedupe=False)
for _ in range(3):
    fn()
ms = [bench.time_gpu(fn, 5, dev) for _ in range(3)]
print(json.dumps({"n_windows": n_win, "layers": L, "gemm_tc": os.environ.get("IQLB_GEMM_TC", "1"), "ms_runs": ms,
                  "windows_per_s": n_win / (min(ms) * 1e-3)}))
