"""Micro-benchmark of iqlb_relabel_mlp: rows/s and roofline fractions, fp32 FFMA path vs split-bf16 tcgen05 path.
Prints one JSON line per case (CUDA events, 3 warm-up + 10 timed calls, inputs resident in HBM)."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from iqlpref_b200.relabel import ensemble_predict, pack_members  # noqa: E402

CASES = [  # name, rows, row stride, d_in, width, depth, members, reduce, alpha
    ("antmaze_q_mlp", 1_000_000, 68, 37, 128, 3, 1, "mean", 0.0),
    ("antmaze_ens10_cvar", 1_000_000, 68, 37, 128, 3, 10, "cvar", 0.9),
    ("pen_bnn40_cvar", 1_000_000, 140, 69, 64, 3, 40, "cvar", 0.9),
    ("pen_human_bnn2440", 5_000, 140, 69, 64, 3, 2440, "cvar", 0.95),
    ("pen_human_ens981", 5_000, 140, 69, 128, 3, 981, "cvar", 0.95),
]


def main():
    dev = torch.device("cuda:0")
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
    except Exception:
        pass
    only = os.environ.get("RELABEL_CASES")
    precs = os.environ.get("RELABEL_PRECISIONS", "split_bf16,fp32").split(",")
    for name, rows, ld, d_in, w, depth, S, reduce, alpha in CASES:
        if only and name not in only.split(","):
            continue
        rng = np.random.default_rng(0)
        members = []
        for _ in range(S):
            m, n_in = [], d_in
            for _l in range(depth):
                m += [rng.standard_normal((n_in, w), dtype=np.float32), rng.standard_normal(w, dtype=np.float32)]
                n_in = w
            m += [rng.standard_normal((w, 1), dtype=np.float32), rng.standard_normal(1, dtype=np.float32)]
            members.append(m)
        packed = pack_members(members, dev)
        table = torch.randn((rows, ld), device=dev)
        x = table[:, :d_in]
        flops = rows * S * 2.0 * (d_in * w + (depth - 1) * w * w + w)
        res = {"case": name, "rows": rows, "members": S, "d_in": d_in, "width": w, "depth": depth, "reduce": reduce}
        for prec in precs:
            if prec == "fp32" and rows * S > 20_000_000:
                n_eff = max(256, 20_000_000 // S)  # bounded sample for the slow path
            else:
                n_eff = rows
            xs = x[:n_eff]
            for _ in range(3):
                ensemble_predict(packed, xs, reduce, alpha, precision=prec, device=dev)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            reps = 10
            e0.record()
            for _ in range(reps):
                ensemble_predict(packed, xs, reduce, alpha, precision=prec, device=dev)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / reps
            res[prec] = {"rows": n_eff, "ms": ms, "rows_per_s": n_eff / (ms * 1e-3),
                         "member_rows_per_s": n_eff * S / (ms * 1e-3),
                         "tflops_fp32_equiv": flops * (n_eff / rows) / (ms * 1e-3) / 1e12,
                         "table_GBps": n_eff * ld * 4 / (ms * 1e-3) / 1e9}
        if "split_bf16" in res and "fp32" in res:
            res["speedup"] = res["split_bf16"]["member_rows_per_s"] / res["fp32"]["member_rows_per_s"]
        print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
