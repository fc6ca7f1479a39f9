#!/usr/bin/env python
"""Per-tile clock64 timeline of CTA 0 of the persistent forward kernel (IQLB_TIMELINE=1).
Needs a library built with the stamps: `make -C iqlpref_b200/csrc clean && make -C iqlpref_b200/csrc TIMELINE=1` (the default
build leaves them out; the per-SM wave tables of the gradient kernels work with either build)."""
import os, sys, argparse
os.environ["IQLB_TIMELINE"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench

ap = argparse.ArgumentParser()
ap.add_argument("--agents-per-gpu", type=int, default=64)
ap.add_argument("--rows", type=int, default=1_000_000)
ap.add_argument("--precision", default="bf16")
args = ap.parse_args()
args.scaling, args.agents_total, args.pt_windows, args.overlap, args.skip_relabel = "weak", 64, 0, 0, True
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
eng, buf, cols, variant, _ = bench.build_workload(args, dev, 0, 1)
for _ in range(5):
    eng.train(None)
torch.cuda.synchronize()
ts = eng.debug_tensor(8).view(torch.int64).cpu().numpy()
def show(name, base, n):
    print(name)
    t0 = ts[0]
    for it in range(8):
        r = ts[base + it * 8: base + it * 8 + n]
        if r[0] == 0:
            break
        print(f"  it{it}: " + " ".join(f"{int(x - t0):7d}" for x in r))
show("epilogue: start, MMA0 ok, epi0 end, BW ok, MMA1 ok, epi1 end, barrier, pre-epi1 barrier", 0, 8)
show("tile end: load_bw issued, stores issued, heads written", 1536, 3)
show("mma: L0 issue, kb0..kb3 issue", 512, 5)
show("producer: X issue, ring e0..e4 issue", 1024, 6)

l = ts[3072:3072 + 8]
if l[0]:
    print("loss CTA 0: " + ", ".join(f"{n}={int(x - l[0])}" for n, x in zip(["start", "loads done", "loop done", "sync", "sums written", "scalars start", "scalars end"], l) if x))

fw = ts[7700:7700 + 296].reshape(148, 2)
if fw[:, 1].any():
    t0 = int(fw[:, 0][fw[:, 0] > 0].min())
    dur = [(int(b_) - t0) for a_, b_ in fw if b_]
    st = [(int(a_) - t0) for a_, b_ in fw if a_]
    import numpy as _np
    print(f"forward CTAs (global timer, ns from the first start): starts {min(st)}..{max(st)}, ends min {min(dur)} median "
          f"{int(_np.median(dur))} max {max(dur)}; ends by CTA index (every 8th): " + " ".join(str(int(fw[i, 1]) - t0) for i in range(0, 148, 8)))
w = ts[3200:3200 + 128]
if w.any():
    print("bwd_weight CTA 0, clocks from the first stamp; per tile: issuer [tile start, P1 operands in, prev dW1 drained, P1 MMAs done, "
          "P2 operands in, H2 in, P2 MMAs done, prev final drain read] | drain [start, dW1 ready, dW1 drained, (end), P2 ready]")
    t0 = min(int(x) for x in w if x)
    for it in range(8):
        r = w[it * 16: it * 16 + 13]
        if not r.any():
            break
        f = lambda x: f"{int(x) - t0:7d}" if x else "      -"
        print(f"  it={it}: i " + " ".join(f(x) for x in r[0:8]) + "  | d " + " ".join(f(x) for x in (r[8], r[9], r[10], r[11], r[12])))
d = ts[2560:2560 + 12]
if os.environ.get("IQLB_BD_PERSIST", "1") != "0":
    print("bwd_data (persistent) CTA 0, clocks from the first stamp; per half-tile n: worker [unit start, half start, loads issued, "
          "A free, produced, drained prev] | issuer [W1+TMEM ready, MMAs issued, MMAs done, dH2 store read]")
    t0 = min(int(x) for x in ts[2560:2560 + 128] if x)
    for n in range(8):
        r = ts[2560 + n * 16: 2560 + n * 16 + 12]
        if not r.any():
            break
        f = lambda x: f"{int(x) - t0:7d}" if x else "      -"
        print(f"  n={n}: w " + " ".join(f(x) for x in r[0:6]) + "  | i " + " ".join(f(x) for x in r[8:12]))
elif d[0]:
    print("bwd_data CTA (0,1,0): " + ", ".join(f"{n}={int(x - d[0])}" for n, x in zip(["start", "loads in regs", "sync", "dH2 tile done", "sync2", "mma0", "mma1", "mma2", "mma3", "drained"], d) if x))

import collections
wv = ts[4096:4096 + 3 * 512].reshape(512, 3)
if wv[:, 1].any():
    by_sm = collections.defaultdict(list)
    for sm, a, b in wv:
        if a:
            by_sm[int(sm)].append((int(a), int(b)))
    t0 = min(a for v in by_sm.values() for a, _ in v)
    print("bwd_data CTAs per SM (globaltimer ns from the first CTA start): [start, end] ...")
    for sm in sorted(by_sm)[:6] + sorted(by_sm)[-2:]:
        print(f"  SM {sm:3d}: " + "  ".join(f"[{a - t0:6d},{b - t0:6d}]" for a, b in sorted(by_sm[sm])))
    dur = [b - a for v in by_sm.values() for a, b in v]
    ends = max(b for v in by_sm.values() for _, b in v)
    gaps = []
    for v in by_sm.values():
        v = sorted(v)
        gaps += [v[i + 1][0] - v[i][1] for i in range(len(v) - 1)]
    import numpy as np
    print(f"  CTA duration ns: median {np.median(dur):.0f}, max {max(dur)}; gap between CTAs on an SM: median {np.median(gaps):.0f};"
          f" kernel span {ends - t0} ns; CTAs per SM: {collections.Counter(len(v) for v in by_sm.values())}")

ww = ts[4096 + 1536:4096 + 1536 + 4 * 512].reshape(512, 4)
if ww[:, 1].any():
    by_sm = collections.defaultdict(list)
    for sm, a, b, net in ww:
        if a:
            by_sm[int(sm)].append((int(a), int(b), int(net)))
    t0 = min(a for v in by_sm.values() for a, _, _ in v)
    print("bwd_weight CTAs per SM: [start, end, net] ...")
    for sm in sorted(by_sm)[:6]:
        print(f"  SM {sm:3d}: " + "  ".join(f"[{a - t0:6d},{b - t0:6d},{n}]" for a, b, n in sorted(by_sm[sm])))
    for net in range(4):
        d = [b - a for v in by_sm.values() for a, b, n in v if n == net]
        print(f"  net {net}: median CTA duration {np.median(d):.0f} ns (max {max(d)})")
    ends = max(b for v in by_sm.values() for _, b, _ in v)
    print(f"  kernel span {ends - t0} ns; CTAs per SM: {collections.Counter(len(v) for v in by_sm.values())}")
