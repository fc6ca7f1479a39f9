"""Micro-benchmark of the MR preference-model training step (iqlb_mr_train_step).
antmaze MR shape (37 -> 128 x 3 -> 1, relu), batch 256 pairs x T-step segments.  Prints one JSON line per case:
device time per step (CUDA events on the launching stream, 5 warm-up + 30 timed steps, inputs resident in HBM, no
per-step host sync), steps/s through `MRTrainer.train` (with the reference's loss.item() sync and H2D of the batch),
the fp32 FLOP rate, and the numpy oracle timed on the host on the same batch (1 step)."""
import ctypes as C
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from iqlpref_b200 import _lib  # noqa: E402
from iqlpref_b200.mr_training import MRTrainer  # noqa: E402
from iqlpref_b200.reward_models.q_mlp import MLP  # noqa: E402
from oracle.mr_train_oracle import MRTrainerOracle, synth_batch  # noqa: E402  (CPU baseline leg only)

CASES_ALL = [("antmaze_T25", 29, 8, [128] * 3, 256, 25), ("antmaze_T100", 29, 8, [128] * 3, 256, 100),
         ("pen_T100_w256", 45, 24, [256] * 3, 256, 100)]
_only = os.environ.get("MR_CASES")
CASES = [c for c in CASES_ALL if not _only or c[0] in _only.split(",")]


def main():
    dev = torch.device("cuda:0")
    for name, S, A, widths, B, T in CASES:
        rng = np.random.default_rng(0)
        torch.manual_seed(0)
        net = MLP(S + A, 1, widths, "relu")
        w0 = [p.detach().numpy().copy() for p in net.parameters()]
        opt = torch.optim.Adam(net.parameters(), lr=3e-4)
        tr = MRTrainer(net, opt, 1, device=dev, precision=os.environ.get("MR_PRECISION", "fp32"))
        batch = synth_batch(rng, B, T, S, A)
        tb = [torch.from_numpy(x) for x in batch]
        for _ in range(3):
            tr.train(tb)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        n_api = 20
        for _ in range(n_api):
            tr.train(tb)
        torch.cuda.synchronize()
        api_ms = (time.perf_counter() - t0) / n_api * 1e3
        # device-resident loop straight on the C ABI
        lib = _lib.load()
        ts = [t.to(dev, torch.float32).contiguous() for t in (tb[0], tb[1], tb[3], tb[4], tb[5], tb[7], tb[8])]
        out2 = torch.zeros(2, device=dev)
        st = _lib.stream_ptr(dev)

        def step(k):
            _lib.check(lib.iqlb_mr_train_step(tr._handle, _lib.ptr(tr.flat), _lib.ptr(tr.exp_avg), _lib.ptr(tr.exp_avg_sq),
                                              *[_lib.ptr(t) for t in ts], B, T, S, A, 3e-4, 0.9, 0.999, 1e-8, k,
                                              _lib.ptr(out2), None, st), "iqlb_mr_train_step")
        k = tr.step_count
        for _ in range(5):
            k += 1; step(k)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        n_dev = 30
        e0.record()
        for _ in range(n_dev):
            k += 1; step(k)
        e1.record()
        torch.cuda.synchronize()
        dev_ms = e0.elapsed_time(e1) / n_dev
        dims = [S + A] + widths
        macs = sum(dims[i] * dims[i + 1] for i in range(len(widths)))
        rows = 2 * B * T
        flop = 2.0 * rows * (macs * 3 - dims[0] * dims[1] + 2 * widths[-1])  # fwd + dW + dX (no dX below layer 0)
        orc = MRTrainerOracle(w0, "relu")
        orc.train(batch)
        t0 = time.perf_counter()
        orc.train(batch)
        cpu_ms = (time.perf_counter() - t0) * 1e3
        print(json.dumps({"case": name, "precision": tr.precision, "rows": rows, "d_in": S + A, "widths": widths, "device_ms_per_step": round(dev_ms, 4),
                          "device_steps_per_s": round(1e3 / dev_ms, 1), "fp32_tflops": round(flop / dev_ms / 1e9, 2),
                          "api_ms_per_step": round(api_ms, 4), "api_steps_per_s": round(1e3 / api_ms, 1),
                          "h2d_bytes_per_step": int(sum(t.numel() * 4 for t in ts)),
                          "cpu_oracle_ms_per_step": round(cpu_ms, 2), "cpu_threads": torch.get_num_threads(),
                          "loss": float(out2[0])}), flush=True)


if __name__ == "__main__":
    main()
