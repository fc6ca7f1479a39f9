"""Micro-benchmark of the preference-transformer training step (iqlb_pt_train_step, SURVEY.md §8f-1, BASELINE configs[2]).
Reference shape (scripts_pt/*.yaml: embd_dim 2^8, head_dim 2^6 -> 4 heads, intermediate 4E, 1 layer, batch 64 pairs, dropouts
0.1) on antmaze-shaped len-100 segment pairs (29 obs, 8 act).  Prints one JSON line per case: device time per step (CUDA
events on the launching stream, 5 warm-up + 20 timed steps, inputs resident in HBM, no per-step host sync), steps/s through
`PTTrainer.train` (H2D of the batch + the reference's loss.item() sync), the algorithmic FLOP rate of the Linear layers, and
the torch oracle timed on the host cores on the same batch (CPU baseline leg; PT_CPU=0 skips it)."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from iqlpref_b200 import _lib  # noqa: E402
from iqlpref_b200.pt_training import PTTrainer  # noqa: E402
from iqlpref_b200.reward_models.pref_transformer import PT  # noqa: E402

# name: (S, A, T, B, E, heads, I, L)
CASES_ALL = {"antmaze_E256_L1_B64": (29, 8, 100, 64, 256, 4, 1024, 1), "antmaze_E256_L1_B256": (29, 8, 100, 256, 256, 4, 1024, 1),
             "pen_E128_L2_B64": (45, 24, 100, 64, 128, 2, 512, 2)}
_only = os.environ.get("PT_CASES")
CASES = {k: v for k, v in CASES_ALL.items() if not _only or k in _only.split(",")}


def main():
    dev = torch.device("cuda:0")
    precision = os.environ.get("PT_PRECISION", "split_bf16")
    for name, (S, A, T, B, E, heads, I, L) in CASES.items():
        from oracle.pt_train_oracle import PTTrainerOracle, synth_pt_batch  # CPU baseline leg + batch synthesis
        rng = np.random.default_rng(0)
        torch.manual_seed(0)
        net = PT(S, A, 1000, E, E, heads, 0.1, 0.1, I, L, 0.1, 2048, 0.1)
        w0 = {k: v.detach().numpy().copy() for k, v in net.state_dict().items() if "causal_bias" not in k}
        opt = torch.optim.Adam(net.parameters(), lr=3e-4)
        tr = PTTrainer(net, opt, device=dev, precision=precision, seed=1)
        batch = synth_pt_batch(rng, B, T, S, A)
        tb = [torch.from_numpy(x) for x in batch]
        for _ in range(3):
            tr.train(tb)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        n_api = 10
        for _ in range(n_api):
            log = tr.train(tb)
        torch.cuda.synchronize()
        api_ms = (time.perf_counter() - t0) / n_api * 1e3
        lib = _lib.load()
        ts, _, _ = tr._unpack(tb)
        out2 = torch.zeros(2, device=dev)
        st = _lib.stream_ptr(dev)

        def step(k):
            _lib.check(lib.iqlb_pt_train_step(tr._handle, _lib.ptr(tr.flat), _lib.ptr(tr.exp_avg), _lib.ptr(tr.exp_avg_sq),
                                              *[_lib.ptr(t) for t in ts], B, 3e-4, 0.9, 0.999, 1e-8, k, 1, _lib.ptr(out2), None,
                                              st), "iqlb_pt_train_step")
        k = tr.step_count
        for _ in range(5):
            k += 1; step(k)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        n_dev = 20
        e0.record()
        for _ in range(n_dev):
            k += 1; step(k)
        e1.record()
        torch.cuda.synchronize()
        dev_ms = e0.elapsed_time(e1) / n_dev
        n, M = 2 * B, 2 * B * 2 * T
        P = E
        lin = M * L * (3 * E * E + E * E + 2 * E * I) + n * T * (S + A) * E + n * T * (2 * P + 1) * E  # MACs of the Linear layers
        attn = n * L * heads * (2 * T) * (2 * T) * (E // heads) + n * T * T * (P + 1)                  # QK^T + PV (causal: half)
        flop = 2.0 * 3 * lin + 2.0 * 3 * attn
        rec = {"case": name, "precision": precision, "pairs": B, "tokens": M, "embd": E, "layers": L,
               "device_ms_per_step": round(dev_ms, 4), "device_steps_per_s": round(1e3 / dev_ms, 2),
               "api_ms_per_step": round(api_ms, 4), "api_steps_per_s": round(1e3 / api_ms, 2),
               "algorithmic_gflop_per_step": round(flop / 1e9, 2), "tflops": round(flop / dev_ms / 1e9, 2),
               "last_loss": log["training_loss"]}
        if os.environ.get("PT_CPU", "1") != "0":
            torch.set_num_threads(os.cpu_count() or 1)
            orc = PTTrainerOracle(w0, heads, 0.1, lr=3e-4)
            t0 = time.perf_counter()
            orc.train(batch)
            rec["cpu_oracle_ms_per_step"] = round((time.perf_counter() - t0) * 1e3, 1)
            rec["cpu_threads"] = torch.get_num_threads()
        print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()
