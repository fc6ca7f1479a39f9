"""Where the host-fed step differs from the sampled one: times, with CUDA events, K steps of
  (a) eng.train(None)              on-device sampling (fused gather + pack)
  (b) eng.train(device_rows)       explicit device batch (D2D copy + pack_x): the host path without PCIe
  (c) eng.train_host_pipelined     pinned host rows every step
and prints one JSON line.  Usage: python scripts/e2e_probe.py [--steps 300]"""
import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=300)
    a = ap.parse_args()
    args = argparse.Namespace(agents_per_gpu=64, rows=1_000_000, precision="bf16", workload="cfg2", scaling="weak",
                              agents_total=64, overlap=0, skip_relabel=True, pt_windows=0, gpus=1)
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    eng, buf, cols, variant, _ = bench.build_workload(args, dev, 0, 1)
    N, B, K = 64, 256, a.steps
    stream = torch.cuda.current_stream(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ring = 4
    dev_rows = []
    host_rows = torch.empty((ring, N, B, eng.R), dtype=torch.float32).pin_memory()
    for i in range(ring):
        idx = torch.randint(0, buf.n_rows, (N, B), device=dev)
        br = buf._table[idx].contiguous()
        dev_rows.append(br)
        host_rows[i].copy_(br)
    host_losses = torch.empty((N, 3), dtype=torch.float32).pin_memory()
    out = {}

    def timed(name, fn, flush=None):
        nonlocal K
        for i in range(20):
            fn(i)
        if flush:
            flush()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e0.record(stream)
        for i in range(K):
            fn(i)
        if flush:
            flush()
        eng.sync()
        e1.record(stream)
        torch.cuda.synchronize()
        out[name] = {"us_per_step_events": round(e0.elapsed_time(e1) * 1e3 / K, 2),
                     "us_per_step_wall": round((time.perf_counter() - t0) * 1e6 / K, 2)}

    for K in (20, 100, 300):  # fixed cost vs per-step cost of the host-fed path
        timed(f"host_pipelined_K{K}", lambda i: eng.train_host_pipelined(host_rows[i % ring], host_losses),
              flush=lambda: eng.train_host_pipelined(None, host_losses))
        out[f"host_pipelined_K{K}"]["total_ms"] = round(out[f"host_pipelined_K{K}"]["us_per_step_wall"] * K / 1e3, 3)
        timed(f"sampled_K{K}", lambda i: eng.train(None))
        out[f"sampled_K{K}"]["total_ms"] = round(out[f"sampled_K{K}"]["us_per_step_wall"] * K / 1e3, 3)
    K = a.steps
    timed("sampled", lambda i: eng.train(None))
    timed("device_rows", lambda i: eng.train(dev_rows[i % ring]))
    timed("host_pipelined", lambda i: eng.train_host_pipelined(host_rows[i % ring], host_losses),
          flush=lambda: eng.train_host_pipelined(None, host_losses))
    timed("sampled_again", lambda i: eng.train(None))
    # host-side cost of one pipelined call when the device is idle between calls (enqueue + wait for the previous step)
    t0 = time.perf_counter()
    for i in range(50):
        eng.train_host_pipelined(host_rows[i % ring], host_losses)
        torch.cuda.synchronize()
    out["host_call_plus_idle_step_us"] = round((time.perf_counter() - t0) * 1e6 / 50, 1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
