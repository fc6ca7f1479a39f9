"""Micro-benchmark of the dataset preparation kernels at BASELINE configs[1] size (1M antmaze-shaped rows):
CUDA events, 3 warm-up + 10 timed calls, inputs resident in HBM; one JSON line."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from iqlpref_b200.dataset_prep import (compute_mean_std_device, modify_reward_device, normalize_states_device,  # noqa: E402
                                       return_reward_range_device)
from oracle import relabel_oracle as ro  # noqa: E402  (CPU baseline leg only)


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    n, S, cap = 1_000_000, 29, 1000
    g = torch.Generator(device="cuda").manual_seed(0)
    r = torch.randn(n, generator=g, device="cuda")
    term = (torch.rand(n, generator=g, device="cuda") < 1e-3).to(torch.uint8)
    table = torch.randn((n, 68), generator=g, device="cuda")
    obs = table[:, :S]
    peaks = {"hbm_gbs": 6650.0}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
    except Exception:
        pass
    out = {"rows": n}
    ms = timed(lambda: return_reward_range_device(r, term, cap))
    out["return_reward_range"] = {"ms": ms, "rows_per_s": n / ms * 1e3, "GBps": n * 9 / ms / 1e6}
    for nr in (2, 6):
        ms = timed(lambda: modify_reward_device(r.clone(), term, "antmaze-large-diverse-v2", nr, cap))
        ms_clone = timed(lambda: r.clone())
        out[f"modify_reward_{nr}"] = {"ms": ms - ms_clone, "rows_per_s": n / (ms - ms_clone) * 1e3}
    ms = timed(lambda: compute_mean_std_device(obs, 1e-3))
    out["compute_mean_std"] = {"ms": ms, "GBps_table": n * 68 * 4 * 2 / ms / 1e6, "hbm_frac": n * 68 * 4 * 2 / ms / 1e6 / peaks["hbm_gbs"]}
    m, s = compute_mean_std_device(obs, 1e-3)
    ms = timed(lambda: normalize_states_device(obs, m, s))
    out["normalize_states"] = {"ms": ms, "GBps_table": n * 68 * 4 * 2 / ms / 1e6}
    # CPU baseline: the oracle port of the reference's Python loop on a bounded sample
    ns = 100_000
    rh, th = r[:ns].cpu().numpy(), term[:ns].cpu().numpy().astype(bool)
    t0 = time.perf_counter()
    ro.modify_reward(rh, th, "antmaze-large-diverse-v2", 6, cap)
    dt = time.perf_counter() - t0
    out["cpu_modify_reward_6"] = {"rows": ns, "s": dt, "rows_per_s": ns / dt, "kind": "port", "cores": 1}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
