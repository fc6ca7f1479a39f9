#!/usr/bin/env python
"""Headline benchmark: IQL updates/sec (B=256, summed over batched agents) on N B200s.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K --warmup W   # CPU arm: the unmodified reference (baseline/_ref)
    python bench.py --gpus N --scaling strong --agents-total 64      # BASELINE configs[4]: 64 agents split over N GPUs

Workload (BASELINE.json configs[1]): synthetic antmaze-large-diverse-shaped
replay data (1M rows, 29 obs, 8 act), batch 256, 2x256 MLPs, Gaussian actor, `--agents-per-gpu`
independent IQL agents (seeds x normalize_reward variants, one reward column per variant) per GPU.
One "step" = one `sample()+train()` of every agent.  Agents never communicate, so ranks share no
data-path collective ("scaling": "weak": agents per GPU fixed; `--scaling strong` fixes the total instead).
The line also carries: `cpu_baseline` (the reference's own iql.py on the host cores, same workload, plus
configs[0] as written under `configs0`), `reference_cuda` (the same stock-PyTorch code on this GPU, 1 agent),
`roofline` of the dominant kernel with SURVEY 8(d)'s algorithmic bytes, and `relabel` (configs[1..3] passes).

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


ACTOR_DROPOUT = None
WORKLOAD = "cfg2"


def set_workload(name):
    """cfg2 = BASELINE.json configs[1] (antmaze-large-diverse-shaped, the headline); cfg4 = the IQL leg of configs[3]
    (pen-human-shaped: 45 obs, 24 act, actor_dropout 0.1, configs/offline/iql/pen/human_v1.yaml)."""
    global S_DIM, A_DIM, HP, ACTOR_DROPOUT, WORKLOAD
    WORKLOAD = name
    if name == "cfg4":
        S_DIM, A_DIM, ACTOR_DROPOUT = 45, 24, 0.1
        HP = dict(HP, beta=3.0, iql_tau=0.8)


def workload_name(agents_per_gpu, rows):
    if WORKLOAD == "cfg4":
        return (f"configs[3] (IQL leg): IQL on synthetic pen-human-shaped data ({rows} rows, {S_DIM} obs, {A_DIM} act), "
                f"batch {BATCH}, 2x{HID} MLPs, Gaussian actor with actor_dropout {ACTOR_DROPOUT}, {agents_per_gpu} "
                f"independent agents per GPU (seeds x {N_VARIANTS} reward columns)")
    return (f"configs[1]: IQL on synthetic antmaze-large-diverse-shaped data ({rows} rows, {S_DIM} obs, {A_DIM} act), "
            f"batch {BATCH}, 2x{HID} MLPs, Gaussian actor, q_mlp-relabelled rewards, {agents_per_gpu} independent agents "
            f"per GPU (seeds x {N_VARIANTS} normalize_reward columns)")


# ------------------------------------------------------------------------------------------------
# algorithmic work per agent-update (restated in DESIGN.md)
# ------------------------------------------------------------------------------------------------
def flops_per_update(S=None, A=None, H=HID, B=BATCH):
    S, A = S or S_DIM, A or A_DIM

    def fwd(i, o):
        return 2 * B * (i * H + H * H + H * o)

    def bwd(i, o):
        return fwd(i, o) + 2 * B * (H * H + H * o)
    f = 2 * fwd(S, 1) + 4 * fwd(S + A, 1) + fwd(S, A)
    b = bwd(S, 1) + 2 * bwd(S + A, 1) + bwd(S, A)
    return f + b


def param_counts(S=None, A=None, H=HID, gaussian=True):
    S, A = S or S_DIM, A or A_DIM
    net = lambda i, o: H * i + H + H * H + H + o * H + o
    pq = 2 * net(S + A, 1)
    p = pq + net(S, 1) + net(S, A) + (A if gaussian else 0)
    return p, pq


def bytes_per_update(S=None, A=None, B=BATCH):
    """fp32 state streamed once: p,m,v read+write (24 B/param), target read+write (8 B/param), batch rows once."""
    S, A = S or S_DIM, A or A_DIM
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
# CPU reference arm.  The UNMODIFIED reference (`ReplayBuffer.sample()` + `ImplicitQLearning.train()` of
# algorithms/offline/iql.py, bf16 autocast as shipped, eager) imported from /root/reference or from its
# git-ignored, shipped copy baseline/_ref (oracle/make_ref.py): `kind: "reference"`.  Its own parallel model is one
# OS process per agent (tr_sweeps/launch.sh:82-92), so the arm runs one single-threaded process per host core, each
# owning its share of the agents and its own replay buffer.  If the reference cannot be imported the oracle port
# (oracle/iql_oracle.py) runs instead and the line says `kind: "port"`.
# ------------------------------------------------------------------------------------------------
def synth_dataset(rows, S, A, seed):
    """The synthetic D4RL-shaped table of SURVEY 8d (cfg1 / cfg2), as the dict `load_d4rl_dataset` takes."""
    import numpy as np
    rng = np.random.default_rng(seed)
    rew = (rng.uniform(size=rows) < 1e-3).astype(np.float32)
    return {"observations": rng.standard_normal((rows, S), dtype=np.float32),
            "actions": rng.uniform(-1, 1, (rows, A)).astype(np.float32),
            "rewards": rew, "next_observations": rng.standard_normal((rows, S), dtype=np.float32),
            "terminals": rew > 0}


def reference_usable():
    try:
        from oracle.ref_import import reference_available
        return bool(reference_available())
    except Exception:  # noqa: BLE001
        return False


def build_reference_agent(ref, S, A, hp, seed, device, deterministic=False):
    """Networks, optimizers and trainer exactly as the reference's train() builds them (iql.py:1467-1515)."""
    import torch
    torch.manual_seed(seed)
    qf, vf = ref.TwinQ(S, A).to(device), ref.ValueFunction(S).to(device)
    actor = (ref.DeterministicPolicy if deterministic else ref.GaussianPolicy)(S, A, 1.0).to(device)
    kw = {"fused": True} if str(device).startswith("cuda") else {}
    v_opt = torch.optim.Adam(vf.parameters(), lr=hp["vf_lr"], **kw)
    q_opt = torch.optim.Adam(qf.parameters(), lr=hp["qf_lr"], **kw)
    a_opt = torch.optim.Adam(actor.parameters(), lr=hp["actor_lr"], **kw)
    return ref.ImplicitQLearning(1.0, actor, a_opt, qf, q_opt, vf, v_opt, iql_tau=hp["iql_tau"], beta=hp["beta"],
                                 max_steps=hp["max_steps"], discount=hp["discount"], tau=hp["tau"], device=str(device))


def _cpu_worker(wid, n_agents, rows, warmup, steps, barrier, out_q, use_ref, S, A, hp, threads):
    sys.stdout = sys.stderr  # the reference prints ("Dataset size: ..."); stdout carries exactly one JSON line
    import torch
    torch.set_num_threads(threads)
    ds = synth_dataset(rows, S, A, 1000 + wid)
    if use_ref:
        from oracle.ref_import import load_reference_iql
        ref = load_reference_iql()
        buf = ref.ReplayBuffer(S, A, rows, "cpu")
        buf.load_d4rl_dataset(ds)
        agents = [build_reference_agent(ref, S, A, hp, wid * 100 + a, "cpu") for a in range(n_agents)]

        def one_step():
            for tr in agents:
                tr.train(buf.sample(BATCH))  # iql.py:1535-1536
    else:
        from oracle.iql_oracle import IQLOracle, OracleHParams, nets_from_reference_modules
        from iqlpref_b200 import iql as biql
        tab = [torch.from_numpy(ds["observations"]), torch.from_numpy(ds["actions"]), torch.from_numpy(ds["rewards"]),
               torch.from_numpy(ds["next_observations"]), torch.from_numpy(ds["terminals"].astype("float32"))]
        agents = []
        for a in range(n_agents):
            torch.manual_seed(wid * 100 + a)
            qf, vf, ac = biql.TwinQ(S, A), biql.ValueFunction(S), biql.GaussianPolicy(S, A, 1.0)
            agents.append(IQLOracle(*nets_from_reference_modules(qf, vf, ac), OracleHParams(**hp), mode="autocast"))

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


def cpu_reference(steps, warmup, total_agents, rows, procs=None, S=S_DIM, A=A_DIM, hp=None, threads=1):
    """One `sample()+train()` of every agent per step, `procs` processes x `threads` torch threads.
    Returns a dict: updates/s, ms per step, cores used, kind, description."""
    import multiprocessing as mp
    hp = hp or HP
    cores = len(os.sched_getaffinity(0))
    procs = procs or max(1, min(cores, total_agents))
    per = max(1, total_agents // procs)
    use_ref = reference_usable()
    ctx = mp.get_context("spawn")
    barrier, q = ctx.Barrier(procs), ctx.Queue()
    ps = [ctx.Process(target=_cpu_worker, args=(w, per, rows, warmup, steps, barrier, q, use_ref, S, A, hp, threads))
          for w in range(procs)]
    for p in ps:
        p.start()
    res = [q.get() for _ in ps]
    for p in ps:
        p.join()
    t0, t1 = min(r[1] for r in res), max(r[2] for r in res)
    n_upd = procs * per * steps
    what = ("the unmodified reference (algorithms/offline/iql.py ReplayBuffer.sample + ImplicitQLearning.train, bf16 "
            "autocast as shipped, eager)" if use_ref else
            "oracle port of iql.py sample()+train() (bf16 autocast as shipped, eager; the reference was not importable)")
    desc = (f"{what}: {procs} process(es) x {threads} thread(s) x {per} agent(s), {steps} steps after {warmup} warm-up, "
            f"{rows}-row replay buffer per process, obs {S} / act {A}; {cores} host cores visible")
    return {"value": n_upd / (t1 - t0), "ms_per_step": (t1 - t0) / steps * 1e3, "cores": procs * threads,
            "kind": "reference" if use_ref else "port", "sample": desc, "agents": procs * per}


def cpu_cfg1_sweep(rows, steps=60, warmup=10):
    """BASELINE.json configs[0] as written: halfcheetah-shaped (17 obs, 6 act), `rows` rows, ONE seed, the reference's
    iql.py on the CPU; torch thread counts swept, best reported (SURVEY 8d 'CPU baseline timing')."""
    cores = len(os.sched_getaffinity(0))
    hp1 = dict(HP, beta=3.0, iql_tau=0.7)  # configs/offline/iql/halfcheetah/medium_v2.yaml
    table, best = [], None
    th = 1
    while th <= cores:
        r = cpu_reference(steps, warmup, 1, rows, procs=1, S=17, A=6, hp=hp1, threads=th)
        table.append({"threads": th, "updates_per_s": r["value"]})
        if best is None or r["value"] > best["value"]:
            best = dict(r, threads=th)
        th *= 2
    return {"value": best["value"], "unit": UNIT, "cores": best["threads"], "kind": best["kind"],
            "workload": f"configs[0]: IQL offline/iql.py on synthetic halfcheetah-medium-v2-shaped data ({rows} x 17 obs, "
                        f"6 act), batch {BATCH}, 2x{HID} MLPs, 1 seed, CPU", "thread_sweep": table}


def common_config(args, world):
    """The `config` object both arms print (identical for the same flags, so the two lines are comparable)."""
    n = agents_per_gpu(args, world)
    return {"workload": workload_name(n, args.rows), "agents_per_gpu": n, "agents_total": n * world,
            "batch_size": BATCH, "rows": args.rows, "obs_dim": S_DIM, "act_dim": A_DIM, "hidden": HID,
            "precision": "bf16" if args.precision == "bf16" else "fp32", "scaling_mode": args.scaling}


def agents_per_gpu(args, world):
    if args.scaling == "strong":
        if args.agents_total % world:
            raise SystemExit(f"--agents-total {args.agents_total} is not divisible by {world} GPUs")
        return args.agents_total // world
    return args.agents_per_gpu


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    r = cpu_reference(args.steps, args.warmup, agents_per_gpu(args, world), args.rows)
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": common_config(args, world),
            "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": r["kind"],
                             "sample": r["sample"]},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------------------
# the CUDA arm
# ------------------------------------------------------------------------------------------------
def time_gpu(fn, reps, dev):
    """Average milliseconds of `fn()` over `reps` back-to-back calls, CUDA events on the current stream, after the
    caller's warm-up; nothing but the calls sits between the two events."""
    import torch
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(dev)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize(dev)
    return e0.elapsed_time(e1) / reps


def mlp_flops_per_row(d_in, w, depth):
    return 2.0 * (d_in * w + (depth - 1) * w * w + w)


def bench_relabel(args, dev, table, R, peaks):
    """Relabel passes of BASELINE configs[1], [2], [3] on synthetic data: rows/s and roofline fractions.  Every entry is
    timed over >= 10 (the single model: 50) repetitions of the library call with pre-packed weights and preallocated
    outputs, CUDA events around the calls only."""
    import numpy as np
    import torch
    from iqlpref_b200.relabel import ensemble_predict, pack_members, pt_predict
    from iqlpref_b200.reward_models.pref_transformer import PT
    from iqlpref_b200.reward_models.q_mlp import MLP as RewardMLP, member_weights
    rows = table.shape[0]
    d_in = S_DIM + A_DIM
    out = {}
    x = table[:, :d_in]  # row-strided view of the replay table: read in place, no copy
    buf = torch.empty(rows, dtype=torch.float32, device=dev)

    def entry(name, fn, reps, n_rows, S, flops_row, bytes_row, note):
        for _ in range(3):
            fn()
        ms = [time_gpu(fn, reps, dev) for _ in range(3)]  # three runs: they must agree
        sec = min(ms) * 1e-3
        fl = n_rows * S * flops_row
        out[name] = {"ms": min(ms), "ms_runs": ms, "reps": reps, "rows": n_rows, "members": S,
                     "rows_per_s": n_rows / sec, "member_rows_per_s": n_rows * S / sec,
                     "flops_per_member_row": flops_row, "tflops_fp32_equiv": fl / sec / 1e12,
                     "tensor_frac_bf16_mma": 3.0 * fl / sec / 1e12 / peaks["tensor_burst"],
                     "hbm_GBps": n_rows * bytes_row / sec / 1e9, "hbm_frac": n_rows * bytes_row / sec / 1e9 / peaks["hbm"],
                     "note": note}

    # configs[1]: q_mlp RewardMLP(37, 1, [128]*3, relu), random init, the whole replay table
    torch.manual_seed(0)
    rm = RewardMLP(d_in, 1, [128, 128, 128], "relu")
    packed1 = pack_members([member_weights(rm)], dev)
    fr = mlp_flops_per_row(d_in, 128, 3)
    entry("mlp", lambda: ensemble_predict(packed1, x, "mean", device=dev, precision="split_bf16", out=buf), 50, rows, 1, fr,
          R * 4 + 4, "single model, split-bf16 tcgen05 (3 bf16 MMAs per product): compute-bound at ~830 flop/B, so the "
          "70 %-of-HBM target of north_star does not apply (DESIGN.md); tensor_frac is against the burst peak")
    out["mlp_fp32"] = {"ms": min(time_gpu(lambda: ensemble_predict(packed1, x, "mean", device=dev, precision="fp32", out=buf),
                                          3, dev) for _ in range(2))}
    r_hat = ensemble_predict(packed1, x, "mean", device=dev, precision="split_bf16").clone()
    # configs[3] shapes: snapshot ensembles with CVaR
    rng = np.random.default_rng(0)

    def members(S, d, w):
        ms = []
        for _ in range(S):
            m, n_in = [], d
            for _l in range(3):
                m += [rng.standard_normal((n_in, w)).astype(np.float32), (rng.standard_normal(w) * 0.1).astype(np.float32)]
                n_in = w
            ms.append(m + [rng.standard_normal((w, 1)).astype(np.float32), np.zeros(1, np.float32)])
        return pack_members(ms, dev)
    p10 = members(10, d_in, 128)
    entry("ens10_cvar", lambda: ensemble_predict(p10, x, "cvar", 0.9, device=dev, out=buf), 10, rows, 10, fr, R * 4 + 4,
          "10 MR snapshots x the whole table, CVaR alpha 0.9 (n_tail 1), selection in the reduction kernel")
    n_pen = 5000
    xp = torch.randn((n_pen, 69), device=dev)
    bufp = torch.empty(n_pen, dtype=torch.float32, device=dev)
    p981 = members(981, 69, 128)
    entry("ens981_pen", lambda: ensemble_predict(p981, xp, "cvar", 0.95, device=dev, out=bufp), 10, n_pen, 981,
          mlp_flops_per_row(69, 128, 3), 69 * 4 + 4, "pen-human: 981 MR snapshots RewardMLP(69,1,[128]*3) x 5000 rows, CVaR 0.95")
    # configs[2]: preference transformer E=256, L=1, q=100 (scripts_pt/antmaze_large_diverse_pt_antmaze_eval.yaml)
    torch.manual_seed(0)
    pt = PT(state_dim=S_DIM, action_dim=A_DIM, max_episode_steps=1000, embd_dim=256, pref_attn_embd_dim=256, num_heads=4,
            intermediate_dim=1024, num_layers=1, max_pos=2048, eps=0.1).to(dev).eval()
    n_win = args.pt_windows
    q = 100
    obs_all = table[:, :S_DIM].contiguous()
    act_all = table[:, S_DIM:d_in].contiguous()
    start = np.arange(n_win, dtype=np.int32) * 7 % (rows - q)
    seq = np.full(n_win, q, dtype=np.int32)
    seq[::10] = 1 + np.arange(len(seq[::10])) % q  # 10 % left-padded windows
    fn_pt = lambda: pt_predict(pt, obs_all, act_all, start, seq, q, device=dev, dedupe=False)
    fn_pt()
    ms = [time_gpu(fn_pt, 3, dev) for _ in range(2)]
    sec = min(ms) * 1e-3
    fl_min, fl_full = 55e6, 384e6  # last-token-only minimum / full forward, per window (E=256, L=1, q=100)
    out["pt"] = {"ms": min(ms), "ms_runs": ms, "distinct_windows": n_win, "windows_per_s": n_win / sec,
                 "flops_per_window_counted": fl_min, "tflops": n_win * fl_min / sec / 1e12,
                 "tensor_frac": n_win * fl_min / sec / 1e12 / peaks["tensor_burst"],
                 "tensor_frac_full_forward_count": n_win * fl_full / sec / 1e12 / peaks["tensor_burst"],
                 "note": "RewardPT(29, 8, E=256, heads 4, I=1024, L=1), query_length 100, every window distinct (with the "
                         "reference's episode-step-as-row-index behaviour, SURVEY F4, a 1M-row dataset has <= 1000 distinct "
                         "windows: its whole relabel is n_windows=1000 of these); includes the host-side window upload"}
    out["rows_per_s"] = out["mlp"]["rows_per_s"]
    return out, r_hat


def bench_trainers(dev):
    """Reward-model training steps of BASELINE configs[2] / [1] on synthetic preference pairs, device-resident
    batches, no per-step host sync: the preference transformer at the reference's shape (scripts_pt/*.yaml: E = 2^8, 4 heads,
    I = 4E, 1 layer, dropouts 0.1, batch 64 pairs x 100 steps) and the MR reward MLP (37 -> 128 x 3 -> 1, 256 pairs x 100)."""
    import numpy as np
    import torch
    from iqlpref_b200.mr_training import MRTrainer
    from iqlpref_b200.pt_training import PTTrainer
    from iqlpref_b200.reward_models.pref_transformer import PT
    from iqlpref_b200.reward_models.q_mlp import MLP as RewardMLP
    out = {}
    rng = np.random.default_rng(0)

    def pairs(B, T, with_ts):
        b = []
        for _ in range(2):
            b += [torch.randn((B, T, S_DIM), device=dev), torch.rand((B, T, A_DIM), device=dev) * 2 - 1,
                  torch.arange(T, device=dev).repeat(B, 1), torch.ones((B, T), device=dev)]
        y = torch.from_numpy((rng.uniform(size=B) < 0.5).astype(np.float32)).to(dev)
        return b + [torch.stack([y, 1 - y], dim=1)]
    torch.manual_seed(0)
    net = PT(S_DIM, A_DIM, 1000, 256, 256, 4, 0.1, 0.1, 1024, 1, 0.1, 2048, 0.1)
    tr = PTTrainer(net, torch.optim.Adam(net.parameters(), lr=3e-4), device=dev, precision="split_bf16", seed=1)
    batch = pairs(64, 100, True)
    fn = lambda: tr.train(batch, sync=False)
    for _ in range(3):
        fn()
    ms = min(time_gpu(fn, 10, dev) for _ in range(2))
    out["pt_train"] = {"ms_per_step": ms, "steps_per_s": 1e3 / ms, "pairs": 64, "segment_len": 100, "precision": "split_bf16",
                       "note": "PTTrainer.train(sync=False): forward + backward + Adam of RewardPT(29, 8, E=256, heads 4, I=1024, "
                               "L=1) with the reference's dropouts, through the Python mirror (input checks included)"}
    torch.manual_seed(0)
    mlp = RewardMLP(S_DIM + A_DIM, 1, [128, 128, 128], "relu")
    mr = MRTrainer(mlp, torch.optim.Adam(mlp.parameters(), lr=3e-4), 1, device=dev, precision="split_bf16")
    mb = pairs(256, 100, False)
    fm = lambda: mr.train(mb, sync=False)
    for _ in range(3):
        fm()
    ms = min(time_gpu(fm, 20, dev) for _ in range(2))
    out["mr_train"] = {"ms_per_step": ms, "steps_per_s": 1e3 / ms, "pairs": 256, "segment_len": 100, "precision": "split_bf16",
                       "note": "MRTrainer.train(sync=False), RewardMLP(37, 1, [128]*3, relu)"}
    # MRTrainerF: the same step + the functional GP prior of an LCF model at 256 measurement points (d = 6 concepts)
    from iqlpref_b200.gp_prior import LCFModel
    from iqlpref_b200.mr_training import MRTrainerF
    torch.manual_seed(0)
    mlp_f = RewardMLP(S_DIM + A_DIM, 1, [128, 128, 128], "relu")

    def feats(X, device):  # the caller's feature map (function_vect): intercept, two inputs and three bounded non-linear columns
        X = X.double()
        return torch.stack([torch.ones_like(X[:, 0]), X[:, 0], X[:, 1], torch.tanh(X[:, 2]), torch.sin(X[:, 0] * X[:, 1]),
                            X[:, -1] ** 2], dim=1)
    prior = LCFModel(np.ones(6, np.float32), feats, dev)
    mrf = MRTrainerF(mlp_f, torch.optim.Adam(mlp_f.parameters(), lr=3e-4), 25600, prior, device=dev, precision="split_bf16")
    xm = torch.randn((256, S_DIM + A_DIM), device=dev)
    ff = lambda: mrf.train(mb, xm, sync=False)
    for _ in range(3):
        ff()
    ms = min(time_gpu(ff, 20, dev) for _ in range(2))
    out["mrf_train"] = {"ms_per_step": ms, "steps_per_s": 1e3 / ms, "pairs": 256, "segment_len": 100, "n_meas": 256, "n_concepts": 6,
                        "precision": "split_bf16",
                        "note": "MRTrainerF.train(sync=False): the MR step with 256 measurement points as extra rows, the d x d "
                                "Woodbury system in fp64 on the device, log p_GP at the updated weights; Phi(X_M) by a torch "
                                "callable on the device (the reference's function_vect is user code)"}
    return out


def build_workload(args, dev, rank, world):
    import torch
    from iqlpref_b200 import iql as biql
    N, rows = agents_per_gpu(args, world), args.rows
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
    # q_mlp relabel of the whole dataset (configs[1]) + the other relabel passes (rank 0 reports them)
    if getattr(args, "skip_relabel", False):  # kernel experiments: only the reward columns are needed
        from iqlpref_b200.relabel import ensemble_predict, pack_members
        from iqlpref_b200.reward_models.q_mlp import MLP as RewardMLP, member_weights
        torch.manual_seed(0)
        rm = RewardMLP(S_DIM + A_DIM, 1, [128, 128, 128], "relu")
        r_hat = ensemble_predict(pack_members([member_weights(rm)], dev), t[:, :S_DIM + A_DIM], "mean", device=dev).clone()
        relabel = {"skipped": True}
    else:
        relabel, r_hat = bench_relabel(args, dev, t, R, load_peaks())
    # reward columns standing in for normalize_reward 0..7 (iql.py:363-401): affine maps of the relabelled reward
    r = r_hat
    span = (r.max() - r.min()).clamp_min(1e-6)
    cols = torch.stack([r, r - 1, r * (1000.0 / span), r * (1000.0 / span) - 1, (r - r.min()) * (1000.0 / span),
                        (r - r.min()) * (1000.0 / span) - 1, (r - r.min() / 1000.0) * (1000.0 / span),
                        (r - r.min() / 1000.0) * (1000.0 / span) - 1]).contiguous()
    variant = (torch.arange(N, device=dev, dtype=torch.int32) % N_VARIANTS).contiguous()
    seeds = [1 + rank * N + a for a in range(N)]
    eng = biql.BatchedIQL(S_DIM, A_DIM, N, hidden_dim=HID, batch_size=BATCH, device=str(dev), precision=args.precision,
                          hparams=[HP] * N, seeds=seeds, actor_dropout=ACTOR_DROPOUT)
    for a in range(N):
        torch.manual_seed(seeds[a])
        qf, vf = biql.TwinQ(S_DIM, A_DIM), biql.ValueFunction(S_DIM)
        ac = biql.GaussianPolicy(S_DIM, A_DIM, 1.0, dropout=ACTOR_DROPOUT)
        eng.adopt(a, qf, vf, ac, bind=False)
    eng.attach_replay(buf, reward_cols=cols, reward_variant=variant)
    eng.set_overlap(bool(args.overlap))  # training-loop mode: the last group's optimizer runs under the next update
    return eng, buf, cols, variant, relabel


def reference_cuda(args, dev, steps=200, warmup=20):
    """The "same box, stock PyTorch" bar: the unmodified reference's sample() + train() on THIS GPU, one
    agent, eager (the reference wraps its nets in torch.compile, iql.py:1523-1527; Inductor is not exercised here),
    fused Adam and three .item() syncs per step as shipped."""
    import torch
    from oracle.ref_import import load_reference_iql
    saved = sys.stdout
    sys.stdout = sys.stderr
    try:
        ref = load_reference_iql()
        buf = ref.ReplayBuffer(S_DIM, A_DIM, args.rows, str(dev))
        buf.load_d4rl_dataset(synth_dataset(args.rows, S_DIM, A_DIM, 0))
        tr = build_reference_agent(ref, S_DIM, A_DIM, HP, 1, str(dev))
        for _ in range(warmup):
            tr.train(buf.sample(BATCH))
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        for _ in range(steps):
            tr.train(buf.sample(BATCH))
        torch.cuda.synchronize(dev)
        dt = time.perf_counter() - t0
    finally:
        sys.stdout = saved
    del buf, tr
    torch.cuda.empty_cache()
    return {"value": steps / dt, "unit": UNIT, "agents": 1, "ms_per_step": dt / steps * 1e3,
            "what": "unmodified reference iql.py (ReplayBuffer.sample + ImplicitQLearning.train, bf16 autocast, "
                    f"Adam(fused=True), eager) on the same B200, 1 agent, {steps} steps after {warmup} warm-up, wall clock "
                    "around the loop (its .item() calls synchronise every step)"}


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
        "bf16_bwd_weight_adam": ("hbm", 24 * p + 8 * pq),
        "head_fwd": ("hbm", 7 * B * H * 4), "head_bwd": ("hbm", 4 * 2 * B * H * 4), "loss": ("hbm", B * row),
        # ALGORITHMIC bytes of the optimizer = p, m, v read + written (24 B/param) and the Polyak
        # target read + written (8 B/target param); nothing else
        "adam_polyak": ("hbm", 24 * p + 8 * pq),
        "adam_polyak_shadow": ("hbm", 24 * p + 8 * pq),
    }
    # bytes this IMPLEMENTATION moves on top of the algorithm (not credited in `frac`): the gradient read (fp32, or
    # bf16 on the tcgen05 path) and, on the tcgen05 path, the refreshed bf16 weight images
    extra = {"adam_polyak": 4 * p, "adam_polyak_shadow": 2 * p + 2 * (p + pq), "bf16_bwd_weight_adam": 2 * (p + pq)}
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
        if k["kernel"] in extra and sec > 0:
            e["design_extra_bytes"] = extra[k["kernel"]] * n_agents
            e["frac_incl_design_extra"] = (amount + extra[k["kernel"]]) * n_agents / sec / 1e9 / peak
        table.append(e)
        if top is None or e["ms_per_step"] > top["ms_per_step"]:
            top = e
    roof = {"bound": top["bound"], "achieved": top["achieved"], "peak": top["peak"], "unit": top["unit"],
            "frac": top["frac"], "traffic": ncu_traffic(top["kernel"]), "kernel": top["kernel"], "peak_source": peaks["source"],
            "traffic_source": "profiles/r02k_bf16_ncu_full.json (dram__bytes_read.sum + dram__bytes_write.sum per launch, "
                              "64 agents)",
            "algorithmic_per_launch": work.get(top["kernel"], ("", 0))[1] * n_agents,
            "algorithmic_definition": "SURVEY 8(d): 24 B/param (p, m, v read + write) + 8 B/target param (Polyak read + "
                                      "write); gradient reads and bf16 weight-image writes are design_extra_bytes",
            "design_extra_bytes": top.get("design_extra_bytes", 0),
            "frac_incl_design_extra": top.get("frac_incl_design_extra")}
    return roof, table


NCU_NAMES = {"adam_polyak_shadow": "adam_polyak_kernel", "adam_polyak": "adam_polyak_kernel", "bf16_fwd": "bf16_fwd_persist_kernel",
             "bf16_bwd_data": "bf16_bwd_data_kernel", "bf16_bwd_weight": "bf16_bwd_weight_kernel", "loss": "loss_kernel",
             "replay_gather_pack": "replay_gather_pack_kernel"}


def ncu_traffic(kernel):
    """DRAM bytes per launch (read + write) of `kernel` from the committed `ncu --set full` capture, or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "r02k_bf16_ncu_full.json")) as f:
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
        r = cpu_reference(args.cpu_steps, 3, agents_per_gpu(args, world), args.rows)
        cpu_base = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]}
        log(f"cpu_baseline: {r['value']:.1f} updates/s on {r['cores']} cores, kind {r['kind']} ({time.perf_counter() - t0:.1f}s)")
        if not args.skip_cfg1:
            t0 = time.perf_counter()
            cpu_base["configs0"] = cpu_cfg1_sweep(args.rows)
            log(f"cpu configs[0]: {cpu_base['configs0']['value']:.1f} updates/s with {cpu_base['configs0']['cores']} thread(s) "
                f"({time.perf_counter() - t0:.1f}s)")
    ref_cuda = None
    if rank == 0 and world == 1 and not args.skip_reference_cuda and reference_usable():
        ref_cuda = reference_cuda(args, dev)
        log(f"reference_cuda: {ref_cuda['value']:.1f} updates/s (stock PyTorch eager, 1 agent)")

    trainers = None
    if rank == 0 and world == 1 and not getattr(args, "skip_relabel", False):
        trainers = bench_trainers(dev)
        log(f"reward-model training: PT {trainers['pt_train']['steps_per_s']:.0f} steps/s, MR {trainers['mr_train']['steps_per_s']:.0f} steps/s")
    eng, buf, cols, variant, relabel = build_workload(args, dev, rank, world)
    N, K, W = agents_per_gpu(args, world), args.steps, args.warmup
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
    eng.sync()  # the optimizer of the last agent group (internal stream) is inside the timed region
    e1.record(stream)
    torch.cuda.synchronize(dev)
    t_host1 = time.perf_counter()
    barrier()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    launches = eng.launches_per_step(True) * K
    # (measured right after the device-timed region, before the extra second of load the clock sampler needs and the
    #  per-kernel pass: after ~1 s of continuous stepping the same step is ~4 % slower - scripts/e2e_probe.py shows the
    #  sampled and the host-fed step drifting together, 227 -> 238 us - and the two numbers should share a thermal state)
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
    eng.sync()
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
    eng.sync()
    e1.record(stream)
    torch.cuda.synchronize(dev)
    t_w1 = time.perf_counter()
    barrier()
    ms_e2e = max_over_ranks(max(e0.elapsed_time(e1), (t_w1 - t_w0) * 1e3))
    assert torch.isfinite(host_losses).all()
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
                "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
                "dtype": "f32" if args.precision == "fp32" else "bf16", "data": "synthetic",
                "config": common_config(args, world),
                "notes": {"sampling": "on-device Philox gather inside the timed region",
                          "l2": f"inputs larger than L2: {N * (3 * eng.Pstride + eng.PQstride) * 4 / 1e6:.0f} MB of "
                                f"agent state + {args.rows * eng.R * 4 / 1e6:.0f} MB replay table per GPU vs 126 MB L2"
                                + ("" if N >= 32 else " (state of this few agents fits L2: the replay table is the "
                                                      "larger-than-L2 input)")},
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
        if trainers is not None:
            line["reward_model_training"] = trainers
        if "mlp" in relabel:
            relabel["roofline"] = {"bound": "tensor", "achieved": 3.0 * relabel["mlp"]["tflops_fp32_equiv"],
                                   "peak": peaks["tensor_burst"], "unit": "TFLOP/s",
                                   "frac": relabel["mlp"]["tensor_frac_bf16_mma"], "hbm_frac": relabel["mlp"]["hbm_frac"],
                                   "note": "single-model relabel, whole iqlb_relabel_mlp_p call (weight-image prep + "
                                           "kernel), timed alone: burst peak; split-bf16 issues three bf16 MMAs per product"}
        if cpu_base is not None:
            line["cpu_baseline"] = cpu_base
        if ref_cuda is not None:
            line["reference_cuda"] = ref_cuda
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
    ap.add_argument("--skip-cfg1", action="store_true", help="skip the configs[0] CPU thread sweep")
    ap.add_argument("--skip-reference-cuda", action="store_true")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: --agents-per-gpu agents on every GPU; strong: --agents-total agents split over the GPUs "
                         "(BASELINE.json configs[4]: 64 agents sharded across 8 GPUs)")
    ap.add_argument("--agents-total", type=int, default=64)
    ap.add_argument("--pt-windows", type=int, default=4096)
    ap.add_argument("--skip-relabel", action="store_true", help="skip the relabel benchmarks (kernel experiments)")
    ap.add_argument("--overlap", type=int, default=int(os.environ.get("IQLB_OVERLAP_ARG", "1")),
                    help="1: iqlb_agents_set_overlap (deferred join of the optimizer stream, synced before the clock stops)")
    ap.add_argument("--workload", default="cfg2", choices=["cfg2", "cfg4"],
                    help="cfg2: antmaze-shaped headline (configs[1]); cfg4: pen-human-shaped IQL leg of configs[3] "
                         "(45 obs / 24 act, actor_dropout 0.1) - relabel benchmarks and the CPU legs are skipped")
    args = ap.parse_args()
    set_workload(args.workload)
    if args.workload == "cfg4":
        args.skip_relabel = args.skip_cpu_baseline = args.skip_reference_cuda = True
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
