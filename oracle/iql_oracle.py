"""CPU oracle for the IQL training step (TEST INFRASTRUCTURE, not product).

A restatement, in plain functional torch on the CPU, of
`/root/reference/algorithms/offline/iql.py`:

* `MLP` / `TwinQ` / `ValueFunction` / `GaussianPolicy` / `DeterministicPolicy` forwards
  (`:417-543`), `asymmetric_l2_loss` (`:404-405`),
* `ImplicitQLearning.train` and `_update_v/_update_q/_update_policy` (`:581-662`),
* `soft_update` (`:127-129`),
* the third-party pieces the reference calls: `torch.optim.Adam` single-tensor update
  (torch `optim/adam.py:347-547`, betas (0.9, 0.999), eps 1e-8, no weight decay) and the
  RECURSIVE `CosineAnnealingLR.get_lr` (torch `optim/lr_scheduler.py:1449-1455`).

Two modes: `fp32` (no autocast; the statement `algorithms/minari/iql.py:466-488` makes of
the same step, and the 1e-3 parity target) and `autocast` (bf16 autocast as shipped,
`iql.py:650`; the timed CPU baseline).

Parity pinning: the reference ships NO tests or golden vectors for this path
(SURVEY.md §4).  This oracle is pinned against outputs of the reference itself, generated
in the authoring container by `oracle/gen_golden.py` and committed under `tests/golden/`
(`tests/test_oracle_golden.py`: per-step losses and final parameters of both modes, small and
full-size nets), and - where the reference can be imported (`/root/reference`, or the
`baseline/_ref` copy) - directly against it on fresh inputs (`tests/test_oracle_vs_reference.py`).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference` leg
may import this module.
"""
import math
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

EXP_ADV_MAX = 100.0  # iql.py:38
LOG_STD_MIN = -20.0  # iql.py:39
LOG_STD_MAX = 2.0  # iql.py:40

Net = List[Tuple[torch.Tensor, torch.Tensor]]  # [(W[out,in], b[out]), ...] like nn.Linear


@dataclass
class OracleHParams:
    discount: float = 0.99
    tau: float = 0.005
    beta: float = 3.0
    iql_tau: float = 0.7
    vf_lr: float = 3e-4
    qf_lr: float = 3e-4
    actor_lr: float = 3e-4
    max_steps: int = 1_000_000
    deterministic: bool = False


def mlp_forward(net: Net, x: torch.Tensor, out_tanh: bool = False,
                dropout_masks: Optional[Sequence[torch.Tensor]] = None) -> torch.Tensor:
    """iql.py:431-449: Linear->ReLU(->Dropout) per hidden layer, Linear(->Tanh) last."""
    n = len(net)
    for i, (w, b) in enumerate(net):
        x = F.linear(x, w, b)
        if i < n - 1:
            x = torch.relu(x)
            if dropout_masks is not None:
                x = x * dropout_masks[i]  # mask already holds 0 or 1/(1-p)
    if out_tanh:
        x = torch.tanh(x)
    return x


def asymmetric_l2_loss(u: torch.Tensor, tau: float) -> torch.Tensor:
    return torch.mean(torch.abs(tau - (u < 0).float()) * u ** 2)


def cosine_lr_next(lr: float, base_lr: float, last_epoch: int, t_max: int, eta_min: float = 0.0) -> float:
    """One step of the recursive CosineAnnealingLR (torch lr_scheduler.py:1449-1455).
    `last_epoch` is the epoch being entered (>= 1)."""
    if (last_epoch - 1 - t_max) % (2 * t_max) == 0:
        return lr + (base_lr - eta_min) * (1 - math.cos(math.pi / t_max)) / 2
    return (1 + math.cos(math.pi * last_epoch / t_max)) / (
        1 + math.cos(math.pi * (last_epoch - 1) / t_max)
    ) * (lr - eta_min) + eta_min


class _Adam:
    """torch.optim.Adam, single-tensor path (optim/adam.py:457,476,529-547)."""

    def __init__(self, params: List[torch.Tensor], lr: float):
        self.params = params
        self.lr = lr
        self.step_count = 0
        self.exp_avg = [torch.zeros_like(p) for p in params]
        self.exp_avg_sq = [torch.zeros_like(p) for p in params]

    @torch.no_grad()
    def step(self, grads: List[torch.Tensor], beta1=0.9, beta2=0.999, eps=1e-8):
        self.step_count += 1
        t = self.step_count
        bc1 = 1 - beta1 ** t
        bc2 = 1 - beta2 ** t
        step_size = self.lr / bc1
        bc2_sqrt = bc2 ** 0.5
        for p, g, m, v in zip(self.params, grads, self.exp_avg, self.exp_avg_sq):
            m.lerp_(g, 1 - beta1)
            v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
            denom = (v.sqrt() / bc2_sqrt).add_(eps)
            p.addcdiv_(m, denom, value=-step_size)


def _flat(net: Net) -> List[torch.Tensor]:
    return [t for wb in net for t in wb]


class IQLOracle:
    """One IQL agent.  Nets are lists of (W, b) float32 CPU tensors (cloned on entry)."""

    def __init__(self, q1: Net, q2: Net, v: Net, actor: Net, log_std: Optional[torch.Tensor],
                 hp: OracleHParams, mode: str = "fp32"):
        assert mode in ("fp32", "autocast")
        cl = lambda net: [(w.detach().clone().float().requires_grad_(True),
                           b.detach().clone().float().requires_grad_(True)) for w, b in net]
        self.q1, self.q2, self.v, self.actor = cl(q1), cl(q2), cl(v), cl(actor)
        self.qt1 = [(w.detach().clone(), b.detach().clone()) for w, b in self.q1]  # iql.py:565
        self.qt2 = [(w.detach().clone(), b.detach().clone()) for w, b in self.q2]
        self.hp = hp
        self.mode = mode
        self.log_std = None
        if not hp.deterministic:
            assert log_std is not None
            self.log_std = log_std.detach().clone().float().requires_grad_(True)
        self.v_params = _flat(self.v)
        self.q_params = _flat(self.q1) + _flat(self.q2)
        # nn.Module.parameters() yields a module's own parameters before its children's, so torch's actor optimizer
        # sees log_std first (iql.py:463-468); Adam treats tensors independently, the order matters only for the
        # checkpoint's state indices (load_checkpoint below)
        self.a_params = ([self.log_std] if self.log_std is not None else []) + _flat(self.actor)
        self.v_opt = _Adam(self.v_params, hp.vf_lr)
        self.q_opt = _Adam(self.q_params, hp.qf_lr)
        self.a_opt = _Adam(self.a_params, hp.actor_lr)
        self.total_it = 0
        self.sched_epoch = 0  # CosineAnnealingLR.last_epoch

    # -- forwards ---------------------------------------------------------------------
    def _policy_bc(self, obs, act, dropout_masks=None):
        mean = mlp_forward(self.actor, obs, out_tanh=True, dropout_masks=dropout_masks)
        if self.hp.deterministic:
            return torch.sum((mean - act) ** 2, dim=1)  # iql.py:629
        std = torch.exp(self.log_std.clamp(LOG_STD_MIN, LOG_STD_MAX))  # iql.py:473
        # Normal.log_prob (torch distributions/normal.py:87-102)
        var = std ** 2
        log_prob = -((act - mean) ** 2) / (2 * var) - std.log() - math.log(math.sqrt(2 * math.pi))
        return -log_prob.sum(-1)

    def step(self, obs, act, rew, next_obs, done, dropout_masks=None) -> Dict[str, float]:
        """iql.py:639-662.  rew/done are [B] (already squeezed)."""
        hp = self.hp
        self.total_it += 1
        ctx = (torch.amp.autocast("cpu", dtype=torch.bfloat16) if self.mode == "autocast"
               else torch.autocast("cpu", enabled=False))
        out = {}
        with ctx:
            with torch.no_grad():
                next_v = mlp_forward(self.v, next_obs).squeeze(-1)
            # _update_v (iql.py:581-593)
            sa = torch.cat([obs, act], 1)
            with torch.no_grad():
                tq = torch.min(mlp_forward(self.qt1, sa).squeeze(-1), mlp_forward(self.qt2, sa).squeeze(-1))
            v = mlp_forward(self.v, obs).squeeze(-1)
            adv = tq - v
            v_loss = asymmetric_l2_loss(adv, hp.iql_tau)
            out["value_loss"] = v_loss.item()
            gv = torch.autograd.grad(v_loss, self.v_params)
            self.v_opt.step(list(gv))
            # _update_q (iql.py:595-613)
            targets = rew + (1.0 - done.float()) * hp.discount * next_v.detach()
            q1 = mlp_forward(self.q1, sa).squeeze(-1)
            q2 = mlp_forward(self.q2, sa).squeeze(-1)
            q_loss = (F.mse_loss(q1, targets) + F.mse_loss(q2, targets)) / 2
            out["q_loss"] = q_loss.item()
            gq = torch.autograd.grad(q_loss, self.q_params)
            self.q_opt.step(list(gq))
            with torch.no_grad():  # soft_update, iql.py:127-129 (uses the UPDATED qf)
                for tnet, snet in ((self.qt1, self.q1), (self.qt2, self.q2)):
                    for (tw, tb), (sw, sb) in zip(tnet, snet):
                        tw.lerp_(sw.detach(), hp.tau)
                        tb.lerp_(sb.detach(), hp.tau)
            # _update_policy (iql.py:615-637)
            exp_adv = torch.exp(hp.beta * adv.detach()).clamp(max=EXP_ADV_MAX)
            bc = self._policy_bc(obs, act, dropout_masks)
            policy_loss = torch.mean(exp_adv * bc)
            out["actor_loss"] = policy_loss.item()
            ga = torch.autograd.grad(policy_loss, self.a_params)
            self.a_opt.step(list(ga))
            self.sched_epoch += 1
            self.a_opt.lr = cosine_lr_next(self.a_opt.lr, hp.actor_lr, self.sched_epoch, hp.max_steps)
        return out

    # -- resume -------------------------------------------------------------------------
    @torch.no_grad()
    def load_checkpoint(self, sd: Dict) -> None:
        """ImplicitQLearning.load_state_dict (iql.py:676-688) on the file the training loop writes
        (`torch.save(trainer.state_dict(), ...)`, :1557-1561): modules, Adam moments and step (state index =
        position in `parameters()`), the scheduler's epoch and the lr it left in the actor optimizer; the target
        is re-created from qf (:679)."""
        mods = (("qf", (("q1.net", self.q1), ("q2.net", self.q2))), ("vf", (("v.net", self.v),)),
                ("actor", (("net.net", self.actor),)))
        for part, nets in mods:
            for prefix, net in nets:
                for i, (w, b) in enumerate(net):
                    w.copy_(sd[part][f"{prefix}.{2 * i}.weight"])
                    b.copy_(sd[part][f"{prefix}.{2 * i}.bias"])
        if self.log_std is not None:
            self.log_std.copy_(sd["actor"]["log_std"])
        for opt, key in ((self.q_opt, "q_optimizer"), (self.v_opt, "v_optimizer"), (self.a_opt, "actor_optimizer")):
            for i in range(len(opt.params)):
                st = sd[key]["state"][i]
                opt.exp_avg[i].copy_(st["exp_avg"])
                opt.exp_avg_sq[i].copy_(st["exp_avg_sq"])
                opt.step_count = int(float(st["step"]))
            opt.lr = float(sd[key]["param_groups"][0]["lr"])
        self.qt1 = [(w.detach().clone(), b.detach().clone()) for w, b in self.q1]
        self.qt2 = [(w.detach().clone(), b.detach().clone()) for w, b in self.q2]
        self.sched_epoch = int(sd["actor_lr_schedule"]["last_epoch"])
        self.total_it = int(sd["total_it"])

    # -- export -------------------------------------------------------------------------
    def export(self) -> Dict[str, np.ndarray]:
        d = {}
        for name, net in (("q1", self.q1), ("q2", self.q2), ("v", self.v), ("actor", self.actor),
                          ("qt1", self.qt1), ("qt2", self.qt2)):
            for i, (w, b) in enumerate(net):
                d[f"{name}.w{i}"] = w.detach().numpy().copy()
                d[f"{name}.b{i}"] = b.detach().numpy().copy()
        if self.log_std is not None:
            d["log_std"] = self.log_std.detach().numpy().copy()
        return d


def nets_from_reference_modules(qf, vf, actor):
    """Extract (q1, q2, v, actor, log_std) lists from reference-shaped nn.Modules."""
    def lin(seq):
        return [(m.weight.detach().clone(), m.bias.detach().clone())
                for m in seq if isinstance(m, torch.nn.Linear)]
    q1 = lin(qf.q1.net)
    q2 = lin(qf.q2.net)
    v = lin(vf.v.net)
    a = lin(actor.net.net)
    log_std = actor.log_std.detach().clone() if hasattr(actor, "log_std") else None
    return q1, q2, v, a, log_std
