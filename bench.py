#!/usr/bin/env python
"""Headline benchmark: IQL updates/sec (B=256, summed over batched agents) on N B200s.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K --warmup W   # CPU reference arm (oracle port)

Workload (BASELINE.json configs[1]): synthetic antmaze-large-diverse-shaped
replay data (1M rows, 29 obs, 8 act), batch 256, 2x256 MLPs, Gaussian actor, `--agents-per-gpu`
independent IQL agents (seeds x normalize_reward variants, one reward column per variant) per GPU.
One "step" = one `sample()+train()` of every agent.  Agents never communicate, so ranks share no
data-path collective ("scaling": "weak": agents per GPU fixed).

One JSON line on stdout (rank 0).  Everything else goes to stderr.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "IQL updates/sec (B=256, summed over batched agents)"
UNIT = "updates/s"
S_DIM, A_DIM, HID, BATCH = 29, 8, 256, 256
HP = dict(discount=0.99, tau=0.005, beta=10.0, iql_tau=0.9, vf_lr=3e-4, qf_lr=3e-4, actor_lr=3e-4,
          max_steps=1_000_000)  # configs/offline/iql/antmaze/large_diverse_v2.yaml
N_VARIANTS = 8


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def workload_name(agents_per_gpu, rows):
    return (f"configs[1]: IQL on synthetic antmaze-large-diverse-shaped data ({rows} rows, {S_DIM} obs, {A_DIM} act), "
            f"batch {BATCH}, 2x{HID} MLPs, Gaussian actor, q_mlp-relabelled rewards, {agents_per_gpu} independent agents "
            f"per GPU (seeds x {N_VARIANTS} normalize_reward columns)")


# ------------------------------------------------------------------------------------------------
# algorithmic work per agent-update (restated in DESIGN.md)
# ------------------------------------------------------------------------------------------------
def flops_per_update(S=S_DIM, A=A_DIM, H=HID, B=BATCH):
    def fwd(i, o):
        return 2 * B * (i * H + H * H + H * o)

    def bwd(i, o):
        return fwd(i, o) + 2 * B * (H * H + H * o)
    f = 2 * fwd(S, 1) + 4 * fwd(S + A, 1) + fwd(S, A)
    b = bwd(S, 1) + 2 * bwd(S + A, 1) + bwd(S, A)
    return f + b


def param_counts(S=S_DIM, A=A_DIM, H=HID, gaussian=True):
    net = lambda i, o: H * i + H + H * H + H + o * H + o
    pq = 2 * net(S + A, 1)
    p = pq + net(S, 1) + net(S, A) + (A if gaussian else 0)
    return p, pq


def bytes_per_update(S=S_DIM, A=A_DIM, B=BATCH):
    """fp32 state streamed once: p,m,v read+write (24 B/param), target read+write (8 B/param), batch rows once."""
    p, pq = param_counts(S, A)
    row = (2 * S + A + 2) * 4
    return 24 * p + 8 * pq + B * row


# ------------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi during the timed region)
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception as e:  # noqa: BLE001
            log("clock sampler unavailable:", e)
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append((time.perf_counter(), ln.strip()))

    def stop(self, t0=None, t1=None):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:  # noqa: BLE001
            self.proc.kill()
        sm, mx, reasons, pw = [], None, set(), []
        for ts, ln in self.lines:
            if t0 is not None and not (t0 - 0.05 <= ts <= t1 + 0.15):
                continue
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1]); pw.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(pw) if pw else None}


# ------------------------------------------------------------------------------------------------
# CPU reference arm: the oracle port of iql.py's sample()+train() (bf16 autocast as shipped), one
# single-threaded process per host core (the reference's own parallel model is one process per agent)
# ------------------------------------------------------------------------------------------------
def _cpu_worker(wid, n_agents, rows, warmup, steps, barrier, out_q):
    import numpy as np
    import torch
    torch.set_num_threads(1)
    from oracle.iql_oracle import IQLOracle, OracleHParams, nets_from_reference_modules
    from iqlpref_b200 import iql as biql
    rng = np.random.default_rng(1000 + wid)
    tab = [torch.from_numpy(rng.standard_normal((rows, S_DIM), dtype=np.float32)),
           torch.from_numpy(rng.uniform(-1, 1, (rows, A_DIM)).astype(np.float32)),
           torch.from_numpy((rng.uniform(size=rows) < 1e-3).astype(np.float32)),
           torch.from_numpy(rng.standard_normal((rows, S_DIM), dtype=np.float32)),
           torch.zeros(rows)]
    agents = []
    for a in range(n_agents):
        torch.manual_seed(wid * 100 + a)
        qf, vf, ac = biql.TwinQ(S_DIM, A_DIM), biql.ValueFunction(S_DIM), biql.GaussianPolicy(S_DIM, A_DIM, 1.0)
        agents.append(IQLOracle(*nets_from_reference_modules(qf, vf, ac), OracleHParams(**HP), mode="autocast"))

    def one_step():
        for ag in agents:
            idx = torch.randint(0, rows, (BATCH,))  # ReplayBuffer.sample, iql.py:211-221
            ag.step(*[t[idx] for t in tab])
    for _ in range(warmup):
        one_step()
    barrier.wait()
    t0 = time.perf_counter()
    for _ in range(steps):
        one_step()
    t1 = time.perf_counter()
    barrier.wait()
    out_q.put((wid, t0, t1))


def cpu_reference(steps, warmup, total_agents, rows=200_000, procs=None):
    """Returns (updates_per_s, ms_per_step, cores_used, description)."""
    import multiprocessing as mp
    cores = len(os.sched_getaffinity(0))
    procs = procs or max(1, min(cores, total_agents))
    per = max(1, total_agents // procs)
    ctx = mp.get_context("spawn")
    barrier, q = ctx.Barrier(procs), ctx.Queue()
    ps = [ctx.Process(target=_cpu_worker, args=(w, per, rows, warmup, steps, barrier, q)) for w in range(procs)]
    for p in ps:
        p.start()
    res = [q.get() for _ in ps]
    for p in ps:
        p.join()
    t0, t1 = min(r[1] for r in res), max(r[2] for r in res)
    n_upd = procs * per * steps
    desc = (f"oracle port of iql.py sample()+train() (bf16 autocast as shipped, eager), {procs} single-thread processes x "
            f"{per} agents, {steps} steps after {warmup} warm-up, {rows}-row tables; {cores} host cores visible")
    return n_upd / (t1 - t0), (t1 - t0) / steps * 1e3, procs, desc, procs * per


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    total = args.agents_per_gpu
    ups, ms, procs, desc, n_agents = cpu_reference(args.steps, args.warmup, total)
    line = {"impl": "reference", "metric": METRIC, "value": ups, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": workload_name(args.agents_per_gpu, args.rows), "agents_in_sample": n_agents,
                       "batch_size": BATCH},
            "cpu_baseline": {"value": ups, "unit": UNIT, "cores": procs, "kind": "port", "sample": desc},
            "e2e": {"value": ups, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------------------
# the CUDA arm
# ------------------------------------------------------------------------------------------------
def build_workload(args, dev, rank):
    import numpy as np
    import torch
    from iqlpref_b200 import iql as biql
    from iqlpref_b200.relabel import ensemble_predict, pack_members
    from iqlpref_b200.reward_models.q_mlp import MLP as RewardMLP, member_weights
    N, rows = args.agents_per_gpu, args.rows
    g = torch.Generator(device=dev)
    g.manual_seed(0)  # the dataset is the same on every rank (replicated)
    buf = biql.ReplayBuffer(S_DIM, A_DIM, rows, device=str(dev))
    R = buf._stride
    t = buf._table
    t[:, :S_DIM] = torch.randn((rows, S_DIM), generator=g, device=dev)
    t[:, S_DIM:S_DIM + A_DIM] = torch.rand((rows, A_DIM), generator=g, device=dev) * 2 - 1
    t[:, S_DIM + A_DIM:2 * S_DIM + A_DIM] = torch.randn((rows, S_DIM), generator=g, device=dev)
    sparse = (torch.rand(rows, generator=g, device=dev) < 1e-3).float()
    t[:, 2 * S_DIM + A_DIM + 1] = sparse  # terminals = goal reached
    buf._size = buf._pointer = rows
    # q_mlp relabel of the whole dataset (configs[1]): RewardMLP(37, 1, [128]*3, relu), random init
    torch.manual_seed(0)
    rm = RewardMLP(S_DIM + A_DIM, 1, [128, 128, 128], "relu")
    x = t[:, :S_DIM + A_DIM]
    packed = pack_members([member_weights(rm)], dev)  # x is a row-strided view of the replay table: no copy
    relabel = {"rows": rows, "model": "RewardMLP(37,1,[128]*3,relu)", "table_row_bytes": R * 4}
    flops = rows * 2.0 * ((S_DIM + A_DIM) * 128 + 2 * 128 * 128 + 128)
    for prec, reps in (("split_bf16", 5), ("fp32", 1)):
        ensemble_predict(packed, x[:4096], "mean", device=dev, precision=prec)  # warm-up
        ensemble_predict(packed, x, "mean", device=dev, precision=prec)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(dev)
        e0.record()
        for _ in range(reps):
            r_hat = ensemble_predict(packed, x, "mean", device=dev, precision=prec)
        e1.record()
        torch.cuda.synchronize(dev)
        ms = e0.elapsed_time(e1) / reps
        relabel[prec] = {"ms": ms, "rows_per_s": rows / (ms * 1e-3), "tflops_fp32_equiv": flops / (ms * 1e-3) / 1e12,
                         "table_GBps": rows * R * 4 / (ms * 1e-3) / 1e9}
    relabel["rows_per_s"] = relabel["split_bf16"]["rows_per_s"]
    relabel["ms"] = relabel["split_bf16"]["ms"]
    # split-bf16 issues three bf16 MMAs per product: the tensor-pipe work is 3x the fp32-equivalent flops
    relabel["bf16_mma_tflops"] = 3.0 * relabel["split_bf16"]["tflops_fp32_equiv"]
    r_hat = ensemble_predict(packed, x, "mean", device=dev, precision="split_bf16")
    # reward columns standing in for normalize_reward 0..7 (iql.py:363-401): affine maps of the relabelled reward
    r = r_hat
    span = (r.max() - r.min()).clamp_min(1e-6)
    cols = torch.stack([r, r - 1, r * (1000.0 / span), r * (1000.0 / span) - 1, (r - r.min()) * (1000.0 / span),
                        (r - r.min()) * (1000.0 / span) - 1, (r - r.min() / 1000.0) * (1000.0 / span),
                        (r - r.min() / 1000.0) * (1000.0 / span) - 1]).contiguous()
    variant = (torch.arange(N, device=dev, dtype=torch.int32) % N_VARIANTS).contiguous()
    seeds = [1 + rank * N + a for a in range(N)]
    eng = biql.BatchedIQL(S_DIM, A_DIM, N, hidden_dim=HID, batch_size=BATCH, device=str(dev), precision=args.precision,
                          hparams=[HP] * N, seeds=seeds)
    for a in range(N):
        torch.manual_seed(seeds[a])
        qf, vf, ac = biql.TwinQ(S_DIM, A_DIM), biql.ValueFunction(S_DIM), biql.GaussianPolicy(S_DIM, A_DIM, 1.0)
        eng.adopt(a, qf, vf, ac, bind=False)
    eng.attach_replay(buf, reward_cols=cols, reward_variant=variant)
    if args.precision == "bf16":
        eng.debug_option(0, int(args.fused_adam))
    return eng, buf, cols, variant, relabel


KERNEL_WORK = None  # filled per precision in kernel_roofline()


def kernel_roofline(prof, n_agents, peaks):
    """Roofline entry for the dominant kernel of the step + the per-kernel table."""
    p, pq = param_counts()
    B, H, S, A = BATCH, HID, S_DIM, A_DIM
    Dq = S + A
    row = (2 * S + A + 2 + 3) // 4 * 4 * 4
    f_l0 = 2 * B * H * (2 * S + 4 * Dq + S)          # 7 first-layer forwards
    f_l1 = 2 * B * H * H * 7
    f_bw1 = 2 * B * H * H * 4
    f_bd1 = 2 * B * H * H * 4
    f_bw0 = 2 * B * H * (S + 2 * Dq + S)
    f_head = 2 * B * H * (6 + A) + 2 * 2 * B * H * (3 + A)
    work = {  # name -> (bound, algorithmic amount per agent: bytes for hbm, flops for tensor)
        "replay_gather": ("hbm", 2 * B * row),
        "replay_gather_pack": ("hbm", 2 * B * row + 2 * B * 128),  # rows read + written, two bf16 images written
        "pack_x": ("hbm", B * row + 2 * B * 128),
        "gemm_fwd_l0": ("tensor", f_l0), "gemm_fwd_l1": ("tensor", f_l1), "gemm_bwd_w1": ("tensor", f_bw1),
        "gemm_bwd_d1": ("tensor", f_bd1), "gemm_bwd_w0": ("tensor", f_bw0),
        "bf16_fwd": ("tensor", f_l0 + f_l1), "bf16_bwd": ("tensor", f_bw1 + f_bd1 + f_bw0),
        "bf16_bwd_data": ("tensor", f_bd1), "bf16_bwd_weight": ("tensor", f_bw1 + f_bw0),
        "bf16_bwd_weight_adam": ("hbm", 24 * p + 8 * pq + 2 * (p + pq)),
        "head_fwd": ("hbm", 7 * B * H * 4), "head_bwd": ("hbm", 4 * 2 * B * H * 4), "loss": ("hbm", B * row),
        "adam_polyak": ("hbm", 24 * p + 8 * pq + 4 * p),  # p,m,v r+w, target r+w, grads read
        "adam_polyak_shadow": ("hbm", 24 * p + 8 * pq + 2 * p + 2 * (p + pq)),  # p,m,v r+w, target r+w, bf16 grads read, bf16 images written
    }
    table, top = [], None
    for k in prof:
        bound, amount = work.get(k["kernel"], ("hbm", 0))
        sec = k["ms_per_step"] * 1e-3
        if bound == "hbm":
            ach, peak, unit = amount * n_agents / sec / 1e9 if sec > 0 else 0.0, peaks["hbm"], "GB/s"
        else:
            ach, peak, unit = amount * n_agents / sec / 1e12 if sec > 0 else 0.0, peaks["tensor"], "TFLOP/s"
        e = {"kernel": k["kernel"], "ms_per_step": k["ms_per_step"], "bound": bound, "achieved": ach, "peak": peak,
             "unit": unit, "frac": ach / peak if peak else None}
        table.append(e)
        if top is None or e["ms_per_step"] > top["ms_per_step"]:
            top = e
    roof = {"bound": top["bound"], "achieved": top["achieved"], "peak": top["peak"], "unit": top["unit"],
            "frac": top["frac"], "traffic": ncu_traffic(top["kernel"]), "kernel": top["kernel"], "peak_source": peaks["source"],
            "traffic_source": "profiles/r01_bf16_ncu_full.json (dram__bytes_read.sum + dram__bytes_write.sum per launch, "
                              "64 agents)",
            "algorithmic_per_launch": work.get(top["kernel"], ("", 0))[1] * n_agents}
    return roof, table


NCU_NAMES = {"adam_polyak_shadow": "adam_polyak_kernel", "adam_polyak": "adam_polyak_kernel", "bf16_fwd": "bf16_fwd_persist_kernel",
             "bf16_bwd_data": "bf16_bwd_data_kernel", "bf16_bwd_weight": "bf16_bwd_weight_kernel", "loss": "loss_kernel",
             "replay_gather_pack": "replay_gather_pack_kernel"}


def ncu_traffic(kernel):
    """DRAM bytes per launch (read + write) of `kernel` from the committed `ncu --set full` capture, or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "r01_bf16_ncu_full.json")) as f:
            prof = json.load(f)
        want = NCU_NAMES.get(kernel, kernel)
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        for k in prof["kernels"]:
            if want in k["kernel"]:
                tot = 0.0
                for key in ("dram_read", "dram_write"):
                    val, unit = k[key].split()
                    tot += float(val) * scale[unit]
                return tot
    except Exception:  # noqa: BLE001
        pass
    return None


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            m = json.load(f)
        return {"hbm": float(m["hbm_gbs"]), "tensor": float(m.get("bf16_tflops_sustained", m["bf16_tflops"])),
                "tensor_burst": float(m["bf16_tflops"]), "source": "MEASURED_PEAKS.json (of measured; bf16 sustained)"}
    except Exception:  # noqa: BLE001
        return {"hbm": 6650.0, "tensor": 1400.0, "tensor_burst": 1590.0, "source": "B200_PROFILING.md fallback (of fallback)"}


def run_cuda(args):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL writes its debug lines (the version banner at NCCL_DEBUG=VERSION) to stdout by default; stdout carries
        # exactly one JSON line
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        # ... and at NCCL_DEBUG=VERSION / WARN it ignores NCCL_DEBUG_FILE: point fd 1 at stderr while the communicator
        # is created (eager with device_id; the barrier covers a lazy one)
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
            os.close(saved_stdout)

    def barrier():
        if world > 1:
            dist.barrier(device_ids=[local])

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    cpu_base = None
    if rank == 0 and world == 1 and not args.skip_cpu_baseline:
        t0 = time.perf_counter()
        ups, ms, procs, desc, _ = cpu_reference(args.cpu_steps, 3, args.agents_per_gpu)
        cpu_base = {"value": ups, "unit": UNIT, "cores": procs, "kind": "port", "sample": desc}
        log(f"cpu_baseline: {ups:.1f} updates/s on {procs} cores ({time.perf_counter() - t0:.1f}s)")

    eng, buf, cols, variant, relabel = build_workload(args, dev, rank)
    N, K, W = args.agents_per_gpu, args.steps, args.warmup
    stream = torch.cuda.current_stream(dev)
    for _ in range(W):
        eng.train(None)
    torch.cuda.synchronize(dev)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_host0 = time.perf_counter()
    e0.record(stream)
    for _ in range(K):
        eng.train(None)
    e1.record(stream)
    torch.cuda.synchronize(dev)
    t_host1 = time.perf_counter()
    barrier()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    launches = eng.launches_per_step(True) * K
    # keep the GPU under the same load while nvidia-smi collects enough samples for a median
    t_extra = time.perf_counter()
    while rank == 0 and (time.perf_counter() - t_host0) < 1.0:
        for _ in range(10):
            eng.train(None)
        torch.cuda.synchronize(dev)
    t_load_end = time.perf_counter() if rank == 0 else t_host1
    del t_extra
    clocks = sampler.stop(t_host0, t_load_end) if rank == 0 else None
    losses = eng.losses.cpu()
    assert torch.isfinite(losses).all(), "non-finite losses in the timed region"

    # per-kernel timing (second pass, same K steps, one event per launch)
    eng.profile(True)
    for _ in range(min(K, 50)):
        eng.train(None)
        eng.profile_collect()
    prof = eng.profile_read()
    eng.profile(False)

    # end to end through the host-buffer entry point: pinned host rows in, losses out, every step
    ring = 4
    host_rows = torch.empty((ring, N, BATCH, eng.R), dtype=torch.float32).pin_memory()
    for i in range(ring):
        rows_i, _ = None, None
        idx = torch.randint(0, buf.n_rows, (N, BATCH), device=dev)
        br = buf._table[idx]  # [N, B, R]
        br[:, :, 2 * S_DIM + A_DIM] = cols[variant.long()[:, None], idx]
        host_rows[i].copy_(br)
    host_losses = torch.empty((N, 3), dtype=torch.float32).pin_memory()
    for i in range(W):
        eng.train_host(host_rows[i % ring], host_losses)
    torch.cuda.synchronize(dev)
    barrier()
    e0.record(stream)
    for i in range(K):
        eng.train_host(host_rows[i % ring], host_losses)
    e1.record(stream)
    torch.cuda.synchronize(dev)
    barrier()
    ms_e2e_sync = max_over_ranks(e0.elapsed_time(e1))
    # pipelined host feed (losses returned with one step of lag; H2D of step k+1 under update k)
    for i in range(W):
        eng.train_host_pipelined(host_rows[i % ring], host_losses)
    eng.train_host_pipelined(None, host_losses)
    torch.cuda.synchronize(dev)
    barrier()
    t_w0 = time.perf_counter()
    e0.record(stream)
    for i in range(K):
        eng.train_host_pipelined(host_rows[i % ring], host_losses)
    eng.train_host_pipelined(None, host_losses)  # flush: losses of the last update are on the host
    e1.record(stream)
    torch.cuda.synchronize(dev)
    t_w1 = time.perf_counter()
    barrier()
    ms_e2e = max_over_ranks(max(e0.elapsed_time(e1), (t_w1 - t_w0) * 1e3))
    assert torch.isfinite(host_losses).all()

    # the path's only training collective: logged loss rows of every rank (outside the timed region)
    from iqlpref_b200 import sharding
    all_losses = sharding.gather_losses(eng.losses, world * N)
    assert tuple(all_losses.shape) == (world * N, 3) and bool(torch.isfinite(all_losses).all())
    if rank == 0:
        peaks = load_peaks()
        roof, table = kernel_roofline(prof, N, peaks)
        ups = world * N * K / (ms_total * 1e-3)
        per_gpu = ups / world
        line = {"metric": METRIC, "value": ups, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
                "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32" if args.precision == "fp32" else "bf16", "data": "synthetic",
                "config": {"workload": workload_name(N, args.rows), "agents_per_gpu": N, "batch_size": BATCH,
                           "precision": args.precision, "sampling": "on-device Philox gather inside the timed region",
                           "l2": f"inputs larger than L2: {N * (3 * eng.Pstride + eng.PQstride) * 4 / 1e6:.0f} MB of "
                                 f"agent state + {args.rows * eng.R * 4 / 1e6:.0f} MB replay table per GPU vs 126 MB L2"},
                "clocks": clocks,
                "e2e": {"value": world * N * K / (ms_e2e * 1e-3), "unit": UNIT,
                        "h2d_bytes_per_step": world * N * BATCH * eng.R * 4, "d2h_bytes_per_step": world * N * 3 * 4,
                        "api": "iqlb_agents_step_host_pipelined (every step: pinned host rows -> H2D on a copy stream -> "
                               "update -> D2H losses, returned with one step of lag; wall clock incl. final flush)",
                        "blocking_api_value": world * N * K / (ms_e2e_sync * 1e-3),
                        "blocking_api": "iqlb_agents_step_host (H2D -> update -> D2H losses -> sync, every step)"},
                "gpu_launches": launches,
                "roofline": roof,
                "kernels": table,
                "step_roofline": {"tensor_frac": flops_per_update() * per_gpu / 1e12 / peaks["tensor"],
                                  "hbm_frac": bytes_per_update() * per_gpu / 1e9 / peaks["hbm"],
                                  "flops_per_update": flops_per_update(), "bytes_per_update": bytes_per_update()},
                "relabel": relabel}
        relabel["roofline"] = {"bound": "tensor", "achieved": relabel["bf16_mma_tflops"], "peak": peaks["tensor_burst"],
                               "unit": "TFLOP/s", "frac": relabel["bf16_mma_tflops"] / peaks["tensor_burst"],
                               "hbm_frac": relabel["split_bf16"]["table_GBps"] / peaks["hbm"],
                               "note": "whole iqlb_relabel_mlp_p call (weight-image prep + kernel), timed alone: burst peak"}
        if cpu_base is not None:
            line["cpu_baseline"] = cpu_base
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--agents-per-gpu", type=int, default=64)
    ap.add_argument("--rows", type=int, default=1_000_000)
    ap.add_argument("--precision", default=os.environ.get("IQLB_PRECISION", "bf16"), choices=["fp32", "bf16"])
    ap.add_argument("--cpu-steps", type=int, default=40)
    ap.add_argument("--skip-cpu-baseline", action="store_true")
    ap.add_argument("--fused-adam", type=int, default=int(os.environ.get("IQLB_FUSED_ADAM", "0")),
                    help="1: Adam/Polyak inside the weight-gradient kernel (gradients never reach HBM)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if args.gpus > 1 and "WORLD_SIZE" not in os.environ:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(29500 + os.getpid() % 1000), os.path.abspath(__file__)] \
            + sys.argv[1:]
        return subprocess.call(cmd)
    return run_cuda(args)


if __name__ == "__main__":
    sys.exit(main())
