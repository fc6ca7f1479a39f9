"""Host-side mirror of `algorithms/offline/iql.py` (reference paths are relative to
/root/reference) for the IQL hot path.  Same names, arguments and errors as the reference;
the arithmetic runs in libiqlb.so (hand-written sm_100a CUDA, include/iqlb.h).

* `TrainConfig`                       iql.py:43-124 (+ `reward_model_root`, iql_eval.py:105-110)
* `ReplayBuffer`                      iql.py:164-226  -> packed table + Philox gather kernel
* `MLP/TwinQ/ValueFunction/GaussianPolicy/DeterministicPolicy`  iql.py:417-543 (parameter
  containers with the reference's state_dict keys; init consumes the torch RNG identically)
* `ImplicitQLearning`                 iql.py:546-688  -> one agent of `BatchedIQL`
* `BatchedIQL`                        new: N independent agents (seeds x sweep hyper-parameters,
                                      tr_sweeps/launch.sh:82-92) advanced by one launch sequence
* `soft_update`, `asymmetric_l2_loss` iql.py:127-129, 404-405

There is no CPU path: constructing any of the compute classes without CUDA + libiqlb.so raises.
"""
import copy
import warnings
import ctypes as C
import math
import os
import uuid
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn as nn
from torch.distributions import Normal

from . import _lib

TensorBatch = List[torch.Tensor]

EXP_ADV_MAX = 100.0
LOG_STD_MIN = -20.0
LOG_STD_MAX = 2.0


@dataclass
class TrainConfig:
    """Field-for-field `TrainConfig` of the reference (iql.py:43-124)."""
    project: str = "IQL-pref"
    group: str = "IQL-D4RL"
    name: str = "IQL"
    env: str = "halfcheetah-medium-expert-v2"
    discount: float = 0.99
    tau: float = 0.005
    beta: float = 3.0
    iql_tau: float = 0.7
    iql_deterministic: bool = False
    max_timesteps: int = int(1e6)
    buffer_size: int = 2_000_000
    batch_size: int = 256
    normalize: bool = True
    normalize_reward: int = 0
    vf_lr: float = 3e-4
    qf_lr: float = 3e-4
    actor_lr: float = 3e-4
    actor_dropout: Optional[float] = None
    log_freq: int = 250
    eval_freq: int = int(5e3)
    n_episodes: int = 10
    checkpoints_path: Optional[str] = None
    load_model: str = ""
    reward_model_path: str = ""
    query_length: int = 1
    bnn_reward_model: bool = False
    bnn_alpha: float = 0.95
    bnn_n_samples: int = 500
    mr_ensemble: bool = False
    mr_alpha: float = 0.95
    mr_burn_in: int = 0
    seed: int = 0
    device: str = "cuda"
    reward_model_root: str = ""  # iql_eval.py:110: reward_model_path = f"{root}_{seed}"

    def __post_init__(self):
        self.name = f"{self.name}-{self.env}-{str(uuid.uuid4())[:8]}"
        if self.checkpoints_path is not None:
            self.checkpoints_path = os.path.join(self.checkpoints_path, self.name)
        if self.reward_model_root and not self.reward_model_path:
            self.reward_model_path = f"{self.reward_model_root}_{self.seed}"


def _require_cuda(device) -> torch.device:
    dev = torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError(f"iqlpref_b200 has no CPU path (device={device!r}); use a CUDA device")
    if not torch.cuda.is_available():
        raise RuntimeError("iqlpref_b200 needs a CUDA device (B200); none is visible")
    if dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    _lib.load()
    return dev


def row_stride(state_dim: int, action_dim: int) -> int:
    """Floats per packed replay row [obs|act|next_obs|rew|done|pad], 16-byte multiple."""
    return (2 * state_dim + action_dim + 2 + 3) // 4 * 4


def pack_rows(obs, act, rew, next_obs, done, stride: int) -> torch.Tensor:
    """Pack the five reference tensors ([n,S],[n,A],[n]|[n,1],[n,S],[n]|[n,1]) into rows."""
    n, s = obs.shape
    a = act.shape[1]
    out = torch.zeros((n, stride), dtype=torch.float32, device=obs.device)
    out[:, :s] = obs
    out[:, s:s + a] = act
    out[:, s + a:2 * s + a] = next_obs
    out[:, 2 * s + a] = rew.reshape(n)
    out[:, 2 * s + a + 1] = done.reshape(n).to(torch.float32)
    return out


def soft_update(target: nn.Module, source: nn.Module, tau: float):
    """iql.py:127-129 with the in-place Polyak kernel (one launch per parameter tensor)."""
    lib = _lib.load()
    for tp, sp in zip(target.parameters(), source.parameters()):
        if not tp.is_cuda:
            raise RuntimeError("soft_update: parameters must live on a CUDA device (no CPU path)")
        with _lib.device_ctx(tp.device):
            if tp.data.is_contiguous() and sp.data.is_contiguous() and tp.data_ptr() % 16 == 0 and sp.data_ptr() % 16 == 0:
                _lib.check(lib.iqlb_polyak(_lib.ptr(tp.data), _lib.ptr(sp.data), tp.numel(), float(tau),
                                           _lib.stream_ptr(tp.device)), "iqlb_polyak")
            else:  # unaligned view of a packed buffer: stage through an aligned copy
                t = tp.data.contiguous().clone()
                s = sp.data.contiguous().clone()
                _lib.check(lib.iqlb_polyak(_lib.ptr(t), _lib.ptr(s), t.numel(), float(tau), _lib.stream_ptr(tp.device)),
                           "iqlb_polyak")
                tp.data.copy_(t)


def asymmetric_l2_loss(u: torch.Tensor, tau: float) -> torch.Tensor:
    """iql.py:404-405 (kept for API parity; the training step computes it in-kernel)."""
    return torch.mean(torch.abs(tau - (u < 0).float()) * u ** 2)


class ReplayBuffer:
    """iql.py:164-226.  Storage is one packed fp32 table [buffer_size, row_stride] in HBM."""

    def __init__(self, state_dim: int, action_dim: int, buffer_size: int, device: str = "cuda"):
        self._device = _require_cuda(device)
        self._state_dim, self._action_dim = state_dim, action_dim
        self._buffer_size = buffer_size
        self._pointer = 0
        self._size = 0
        self._stride = row_stride(state_dim, action_dim)
        self._table = torch.zeros((buffer_size, self._stride), dtype=torch.float32, device=self._device)

    # views with the reference's attribute names
    @property
    def _states(self):
        return self._table[:, :self._state_dim]

    @property
    def _actions(self):
        return self._table[:, self._state_dim:self._state_dim + self._action_dim]

    @property
    def _next_states(self):
        s, a = self._state_dim, self._action_dim
        return self._table[:, s + a:2 * s + a]

    @property
    def _rewards(self):
        c = 2 * self._state_dim + self._action_dim
        return self._table[:, c:c + 1]

    @property
    def _dones(self):
        c = 2 * self._state_dim + self._action_dim + 1
        return self._table[:, c:c + 1]

    def _to_tensor(self, data: np.ndarray) -> torch.Tensor:
        return torch.tensor(data, dtype=torch.float32, device=self._device)

    def load_d4rl_dataset(self, data: Dict[str, np.ndarray]):
        if self._size != 0:
            raise ValueError("Trying to load data into non-empty replay buffer")
        n = data["observations"].shape[0]
        if n > self._buffer_size:
            raise ValueError("Replay buffer is smaller than the dataset you are trying to load!")
        self._states[:n] = self._to_tensor(data["observations"])
        self._actions[:n] = self._to_tensor(data["actions"])
        self._rewards[:n] = self._to_tensor(data["rewards"][..., None])
        self._next_states[:n] = self._to_tensor(data["next_observations"])
        self._dones[:n] = self._to_tensor(data["terminals"][..., None])
        self._size += n
        self._pointer = min(self._size, n)
        print(f"Dataset size: {n}")

    @property
    def n_rows(self) -> int:
        return min(self._size, self._pointer)

    def split(self, rows: torch.Tensor) -> TensorBatch:
        s, a = self._state_dim, self._action_dim
        return [rows[:, :s], rows[:, s:s + a], rows[:, 2 * s + a:2 * s + a + 1], rows[:, s + a:2 * s + a],
                rows[:, 2 * s + a + 1:2 * s + a + 2]]

    def sample_rows(self, batch_size: int, return_indices: bool = False):
        """Packed rows [batch, row_stride].  Draws from torch's CUDA generator state (seed, philox
        offset) and advances it exactly like `torch.randint(0, n, (batch,), device=cuda)`."""
        n = self.n_rows
        if n <= 0:
            raise RuntimeError("cannot sample from an empty replay buffer")
        gen = torch.cuda.default_generators[self._device.index]
        seed, offset = gen.initial_seed(), gen.get_offset()
        rng = torch.tensor([[seed, offset]], dtype=torch.int64, device=self._device)
        out = torch.empty((batch_size, self._stride), dtype=torch.float32, device=self._device)
        idx = torch.empty((1, batch_size), dtype=torch.int64, device=self._device) if return_indices else None
        lib = _lib.load()
        with _lib.device_ctx(self._device):
            _lib.check(lib.iqlb_replay_sample(_lib.ptr(self._table), n, self._stride,
                                              2 * self._state_dim + self._action_dim, None, None, _lib.ptr(rng), 1,
                                              batch_size, _lib.ptr(idx), _lib.ptr(out), _lib.stream_ptr(self._device)),
                       "iqlb_replay_sample")
        gen.set_offset(offset + philox_offset_increment(batch_size, n, self._device))
        return (out, idx[0]) if return_indices else out

    def sample(self, batch_size: int) -> TensorBatch:
        return self.split(self.sample_rows(batch_size))

    def add_transition(self):
        raise NotImplementedError


def philox_offset_increment(batch: int, n: int, device) -> int:
    """Generator advance of one torch.randint call (ATen DistributionTemplates.h:50-63)."""
    props = torch.cuda.get_device_properties(device)
    block = 256
    grid = min(props.multi_processor_count * (props.max_threads_per_multi_processor // block),
               (batch + block - 1) // block)
    unroll = 2 if n >= (1 << 28) else 4
    return ((batch - 1) // (block * grid * unroll) + 1) * 4


# --------------------------------------------------------------------------------------------
# Parameter containers (same module tree / state_dict keys as the reference)
# --------------------------------------------------------------------------------------------
class Squeeze(nn.Module):
    def __init__(self, dim=-1):
        super().__init__()
        self.dim = dim

    def forward(self, x):
        return x.squeeze(dim=self.dim)


class MLP(nn.Module):
    def __init__(self, dims, activation_fn=nn.ReLU, output_activation_fn=None, squeeze_output=False, dropout=None):
        super().__init__()
        if len(dims) < 2:
            raise ValueError("MLP requires at least two dims (input and output)")
        mods: List[nn.Module] = []
        for d_in, d_out in zip(dims[:-2], dims[1:-1]):
            mods += [nn.Linear(d_in, d_out), activation_fn()]
            if dropout is not None:
                mods.append(nn.Dropout(dropout))
        mods.append(nn.Linear(dims[-2], dims[-1]))
        if output_activation_fn is not None:
            mods.append(output_activation_fn())
        if squeeze_output:
            if dims[-1] != 1:
                raise ValueError("Last dim must be 1 when squeezing")
            mods.append(Squeeze(-1))
        self.net = nn.Sequential(*mods)

    def forward(self, x):
        return self.net(x)

    def linears(self) -> List[nn.Linear]:
        return [m for m in self.net if isinstance(m, nn.Linear)]


class GaussianPolicy(nn.Module):
    def __init__(self, state_dim, act_dim, max_action, hidden_dim=256, n_hidden=2, dropout=None):
        super().__init__()
        self.net = MLP([state_dim, *([hidden_dim] * n_hidden), act_dim], output_activation_fn=nn.Tanh, dropout=dropout)
        self.log_std = nn.Parameter(torch.zeros(act_dim, dtype=torch.float32))
        self.max_action = max_action

    def forward(self, obs):
        return Normal(self.net(obs), torch.exp(self.log_std.clamp(LOG_STD_MIN, LOG_STD_MAX)))

    @torch.inference_mode()
    def act(self, state: np.ndarray, device: str = "cpu"):
        state = torch.tensor(state.reshape(1, -1), device=device, dtype=torch.float32)
        dist = self(state)
        action = dist.mean if not self.training else dist.sample()
        return torch.clamp(self.max_action * action, -self.max_action, self.max_action).cpu().data.numpy().flatten()


class DeterministicPolicy(nn.Module):
    def __init__(self, state_dim, act_dim, max_action, hidden_dim=256, n_hidden=2, dropout=None):
        super().__init__()
        self.net = MLP([state_dim, *([hidden_dim] * n_hidden), act_dim], output_activation_fn=nn.Tanh, dropout=dropout)
        self.max_action = max_action

    def forward(self, obs):
        return self.net(obs)

    @torch.inference_mode()
    def act(self, state: np.ndarray, device: str = "cpu"):
        state = torch.tensor(state.reshape(1, -1), device=device, dtype=torch.float32)
        return torch.clamp(self(state) * self.max_action, -self.max_action, self.max_action).cpu().data.numpy().flatten()


class TwinQ(nn.Module):
    def __init__(self, state_dim, action_dim, hidden_dim=256, n_hidden=2):
        super().__init__()
        dims = [state_dim + action_dim, *([hidden_dim] * n_hidden), 1]
        self.q1 = MLP(dims, squeeze_output=True)
        self.q2 = MLP(dims, squeeze_output=True)

    def both(self, state, action):
        sa = torch.cat([state, action], 1)
        return self.q1(sa), self.q2(sa)

    def forward(self, state, action):
        return torch.min(*self.both(state, action))


class ValueFunction(nn.Module):
    def __init__(self, state_dim, hidden_dim=256, n_hidden=2):
        super().__init__()
        self.v = MLP([state_dim, *([hidden_dim] * n_hidden), 1], squeeze_output=True)

    def forward(self, state):
        return self.v(state)


def _mlp_linears(m: nn.Module) -> List[nn.Linear]:
    return [x for x in m.net if isinstance(x, nn.Linear)]


def _check_two_hidden(lins: Sequence[nn.Linear], what: str):
    if len(lins) != 3:
        raise ValueError(f"{what}: the B200 step is built for n_hidden=2 (3 Linear layers), got {len(lins)}")


# --------------------------------------------------------------------------------------------
# N independent agents
# --------------------------------------------------------------------------------------------
def _strip_compile_prefix_key(key: str) -> str:
    """`torch.compile`d modules (iql.py:1523-1527) prefix their state_dict keys with `_orig_mod.`."""
    return key[len("_orig_mod."):] if key.startswith("_orig_mod.") else key


def _to_cpu(obj):
    if isinstance(obj, torch.Tensor):
        return obj.detach().cpu()
    if isinstance(obj, dict):
        return {k: _to_cpu(v) for k, v in obj.items()}
    if isinstance(obj, list):
        return [_to_cpu(v) for v in obj]
    return obj


class BatchedIQL:
    """N independent IQL agents (same network shapes, per-agent hyper-parameters, seeds and
    reward labels) advanced together.  Agent i obeys the contract of the reference's
    `ImplicitQLearning` (iql.py:546-688)."""

    HP_KEYS = ("discount", "tau", "beta", "iql_tau", "vf_lr", "qf_lr", "actor_lr", "max_steps")

    def __init__(self, state_dim: int, action_dim: int, n_agents: int, hidden_dim: int = 256, batch_size: int = 256,
                 deterministic: bool = False, device: str = "cuda", precision: str = "fp32",
                 hparams: Optional[Sequence[Dict[str, float]]] = None, seeds: Optional[Sequence[int]] = None,
                 actor_dropout: Optional[float] = None):
        self.device = _require_cuda(device)
        self.lib = _lib.load()
        self.S, self.A, self.H, self.B, self.N = state_dim, action_dim, hidden_dim, batch_size, n_agents
        self.deterministic = bool(deterministic)
        self.precision = {"fp32": 0, "bf16": 1}[precision]
        self.R = row_stride(state_dim, action_dim)
        p_all, p_q = C.c_int64(), C.c_int64()
        _lib.check(self.lib.iqlb_param_count(state_dim, action_dim, hidden_dim, int(self.deterministic),
                                             C.byref(p_all), C.byref(p_q)), "iqlb_param_count")
        self.P, self.PQ = p_all.value, p_q.value
        self.Pstride, self.PQstride = (self.P + 7) // 8 * 8, (self.PQ + 7) // 8 * 8
        z = lambda *shape, dt=torch.float32: torch.zeros(shape, dtype=dt, device=self.device)
        self.params, self.exp_avg, self.exp_avg_sq = z(n_agents, self.Pstride), z(n_agents, self.Pstride), z(n_agents, self.Pstride)
        self.target = z(n_agents, self.PQstride)
        self.hparams = z(n_agents, _lib.HP_STRIDE)
        self.it = z(n_agents, dt=torch.int64)
        self.actor_lr = z(n_agents, dt=torch.float64)
        self.rng = z(n_agents, 2, dt=torch.int64)
        self.losses, self.loss_sum = z(n_agents, 3), z(n_agents, 3)
        defaults = dict(discount=0.99, tau=0.005, beta=3.0, iql_tau=0.7, vf_lr=3e-4, qf_lr=3e-4, actor_lr=3e-4,
                        max_steps=1_000_000)
        hp_rows, lr0 = [], []
        for i in range(n_agents):
            hp = dict(defaults)
            if hparams is not None:
                hp.update(hparams[i])
            hp_rows.append([float(hp[k]) for k in self.HP_KEYS])
            lr0.append(float(hp["actor_lr"]))
        self.hp_host = hp_rows
        self.hparams.copy_(torch.tensor(hp_rows, dtype=torch.float32))
        self.actor_lr.copy_(torch.tensor(lr0, dtype=torch.float64))
        if seeds is not None:
            self.rng[:, 0] = torch.tensor([int(s) for s in seeds], dtype=torch.int64)
        desc = _lib.AgentsDesc(n_agents, state_dim, action_dim, hidden_dim, batch_size, int(self.deterministic),
                               self.R, self.precision)
        handle = C.c_void_p()
        with _lib.device_ctx(self.device):  # the library allocates its scratch on the current device
            _lib.check(self.lib.iqlb_agents_create(C.byref(desc), _lib.ptr(self.params), _lib.ptr(self.exp_avg),
                                                   _lib.ptr(self.exp_avg_sq), _lib.ptr(self.target), _lib.ptr(self.hparams),
                                                   _lib.ptr(self.it), _lib.ptr(self.actor_lr), _lib.ptr(self.rng),
                                                   _lib.ptr(self.losses), _lib.ptr(self.loss_sum), C.byref(handle)),
                       "iqlb_agents_create")
        self._h = handle
        self._replay = None
        self._layout = self._make_layout()
        self.actor_dropout = float(actor_dropout or 0.0)
        # the reference's MLP inserts nn.Dropout after every hidden ReLU whenever `dropout is not None` (iql.py:436-437),
        # which moves the actor's Linear layers to net.net.0 / 3 / 6 in its state_dict (adopt() re-derives this from
        # the module it is given)
        self._actor_lin_idx = (0, 3, 6) if actor_dropout is not None else (0, 2, 4)
        if self.actor_dropout > 0.0:
            with _lib.device_ctx(self.device):
                _lib.check(self.lib.iqlb_agents_set_actor_dropout(self._h, C.c_float(self.actor_dropout)),
                           "iqlb_agents_set_actor_dropout")

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and getattr(self, "lib", None) is not None:
            self.lib.iqlb_agents_destroy(h)
            self._h = None

    # ---- packed layout (mirrors Layout in csrc/iql_step.cuh) -------------------------------
    def _make_layout(self):
        S, A, H = self.S, self.A, self.H
        lay, off = {}, 0
        for net, d_in, d_out in (("q1", S + A, 1), ("q2", S + A, 1), ("v", S, 1), ("actor", S, A)):
            for name, shape in (("w0", (H, d_in)), ("b0", (H,)), ("w1", (H, H)), ("b1", (H,)), ("w2", (d_out, H)),
                                ("b2", (d_out,))):
                n = int(np.prod(shape))
                lay[f"{net}.{name}"] = (off, shape)
                off += n
            off = (off + 7) // 8 * 8  # every network segment is padded to 8 floats (Layout::init)
        if not self.deterministic:
            lay["log_std"] = (off, (A,))
            off += A
        assert off == self.P, (off, self.P)
        return lay

    def view(self, buf: torch.Tensor, agent: int, key: str) -> torch.Tensor:
        off, shape = self._layout[key]
        n = int(np.prod(shape))
        return buf[agent, off:off + n].view(*shape)

    # ---- adopting reference-shaped modules ----------------------------------------------------
    def _module_tensors(self, qf, vf, actor):
        out = {}
        for net, mod in (("q1", qf.q1), ("q2", qf.q2), ("v", vf.v), ("actor", actor.net)):
            lins = _mlp_linears(mod)
            _check_two_hidden(lins, net)
            for i, lin in enumerate(lins):
                out[f"{net}.w{i}"] = lin.weight
                out[f"{net}.b{i}"] = lin.bias
        if not self.deterministic:
            out["log_std"] = actor.log_std
        return out

    def adopt(self, agent: int, qf: nn.Module, vf: nn.Module, actor: nn.Module, bind: bool = True):
        """Copy the modules' parameters into agent slot `agent` (target := qf, iql.py:565) and, if
        `bind`, re-point the module parameters at the packed storage so `module.state_dict()`
        always reflects the trained values."""
        self.sync()
        for m in list(actor.modules()):
            if isinstance(m, nn.Dropout) and m.p > 0 and abs(m.p - self.actor_dropout) > 1e-12:
                raise ValueError(f"the actor uses Dropout({m.p}) but the engine was built with actor_dropout="
                                 f"{self.actor_dropout}: pass actor_dropout={m.p}")
        lin_idx = tuple(i for i, m in enumerate(actor.net.net) if isinstance(m, nn.Linear))
        if len(lin_idx) == 3:
            self._actor_lin_idx = lin_idx  # state_dict names of the actor's Linear layers (0/2/4, or 0/3/6 with Dropout)
        with torch.no_grad():
            for key, prm in self._module_tensors(qf, vf, actor).items():
                dst = self.view(self.params, agent, key)
                if tuple(prm.shape) != tuple(dst.shape):
                    raise ValueError(f"{key}: module shape {tuple(prm.shape)} != packed shape {tuple(dst.shape)}")
                dst.copy_(prm.detach().to(self.device, torch.float32))
                if bind:
                    prm.data = dst
            self.target[agent, :self.PQ] = self.params[agent, :self.PQ]
        self._sync_shadow()

    @_lib.on_device
    def _sync_shadow(self):
        """params/target were written from Python: rebuild the bf16 operand images (precision bf16)."""
        _lib.check(self.lib.iqlb_agents_refresh(self._h, _lib.stream_ptr(self.device)), "iqlb_agents_refresh")

    # ---- stepping ---------------------------------------------------------------------------------
    @_lib.on_device
    def attach_replay(self, buffer: ReplayBuffer, reward_cols: Optional[torch.Tensor] = None,
                      reward_variant: Optional[torch.Tensor] = None, record_indices: bool = False):
        if buffer._stride != self.R:
            raise ValueError("replay buffer row layout does not match the agents")
        self._replay = buffer
        self._reward_cols, self._reward_variant = reward_cols, reward_variant
        self.last_indices = (torch.empty((self.N, self.B), dtype=torch.int64, device=self.device)
                             if record_indices else None)
        _lib.check(self.lib.iqlb_agents_set_replay(self._h, _lib.ptr(buffer._table), buffer.n_rows,
                                                   _lib.ptr(reward_cols), _lib.ptr(reward_variant),
                                                   _lib.ptr(self.last_indices)), "iqlb_agents_set_replay")

    @_lib.on_device
    def train(self, batch_rows: Optional[torch.Tensor] = None):
        """Enqueue one update of every agent.  `batch_rows`: packed [N, B, row_stride] device rows,
        or None to sample on device.  Losses land in `self.losses` ([N,3]) without a host sync."""
        if batch_rows is not None:
            if tuple(batch_rows.shape) != (self.N, self.B, self.R) or batch_rows.dtype != torch.float32:
                raise ValueError(f"batch_rows must be float32 [{self.N},{self.B},{self.R}]")
            batch_rows = batch_rows.contiguous()
        _lib.check(self.lib.iqlb_agents_step(self._h, _lib.ptr(batch_rows), _lib.stream_ptr(self.device)),
                   "iqlb_agents_step")

    @_lib.on_device
    @torch.no_grad()
    def act(self, states, max_action: float = 1.0) -> torch.Tensor:
        """Deterministic actions of every agent's policy in eval mode (GaussianPolicy.act / DeterministicPolicy.act,
        iql.py:476-482, 506-514; the device half of eval_actor :304-319): clamp(max_action * tanh(actor(s))).
        `states`: [n, state_dim] (all agents evaluate the same rows) or [n_agents, n, state_dim]; returns
        float32 [n_agents, n, action_dim] on the device, fp32 arithmetic like the reference's evaluation."""
        x = torch.as_tensor(states, dtype=torch.float32).to(self.device).contiguous()
        if x.ndim == 1:
            x = x.reshape(1, -1)
        if x.shape[-1] != self.S or x.ndim not in (2, 3) or (x.ndim == 3 and x.shape[0] != self.N):
            raise ValueError(f"states must be [n, {self.S}] or [{self.N}, n, {self.S}]")
        n = x.shape[-2]
        out = torch.empty((self.N, n, self.A), dtype=torch.float32, device=self.device)
        if n:
            stride = 0 if x.ndim == 2 else n * self.S
            _lib.check(self.lib.iqlb_agents_act(self._h, _lib.ptr(x), stride, n, float(max_action), _lib.ptr(out),
                                                _lib.stream_ptr(self.device)), "iqlb_agents_act")
        return out

    @_lib.on_device
    def train_host(self, host_rows: torch.Tensor, host_losses: torch.Tensor):
        """Step with the batch in pinned host memory; returns after the losses are on the host."""
        _lib.check(self.lib.iqlb_agents_step_host(self._h, _lib.ptr(host_rows), _lib.ptr(host_losses),
                                                  _lib.stream_ptr(self.device)), "iqlb_agents_step_host")

    @_lib.on_device
    def train_host_pipelined(self, host_rows: Optional[torch.Tensor], host_losses_prev: torch.Tensor):
        """Pipelined host feed: enqueue the update on `host_rows` (pinned; None = flush) and return the losses of
        the PREVIOUS update in `host_losses_prev`; the H2D copy of the next call overlaps this update."""
        _lib.check(self.lib.iqlb_agents_step_host_pipelined(self._h, _lib.ptr(host_rows) if host_rows is not None else None,
                                                            _lib.ptr(host_losses_prev), _lib.stream_ptr(self.device)),
                   "iqlb_agents_step_host_pipelined")

    @_lib.on_device
    def sync(self):
        """Make the optimizer work the engine runs on its internal stream visible to the current stream (needed only
        after `set_overlap(True)`, before touching `params / exp_avg / exp_avg_sq / target`; see `iqlb_agents_sync`)."""
        _lib.check(self.lib.iqlb_agents_sync(self._h, _lib.stream_ptr(self.device)), "iqlb_agents_sync")

    @_lib.on_device
    def set_overlap(self, enable: bool):
        """Training-loop mode (precision "bf16"): the optimizer of the last agent group keeps running under the first
        group's NEXT update instead of being joined at the end of every `train()` call.  Losses are unaffected; call
        `sync()` before reading parameters or optimizer state.  Checkpoint / export helpers of this class do."""
        _lib.check(self.lib.iqlb_agents_set_overlap(self._h, int(bool(enable)), _lib.stream_ptr(self.device)),
                   "iqlb_agents_set_overlap")

    def launches_per_step(self, sampled: bool) -> int:
        return int(self.lib.iqlb_agents_launches_per_step(self._h, int(sampled)))

    # ---- per-kernel timing (CUDA events on the launching stream; used by bench.py) ----------------------
    @_lib.on_device
    def profile(self, enable: bool):
        _lib.check(self.lib.iqlb_agents_profile(self._h, int(bool(enable))), "iqlb_agents_profile")

    @_lib.on_device
    def profile_collect(self):
        _lib.check(self.lib.iqlb_agents_profile_collect(self._h), "iqlb_agents_profile_collect")

    @_lib.on_device
    def profile_read(self) -> List[Dict[str, Any]]:
        """[{"kernel": name, "ms_per_step": t}] for every launch of one step, in launch order."""
        out, i = [], 0
        name, ms, steps = C.c_char_p(), C.c_double(), C.c_int32()
        while self.lib.iqlb_agents_profile_read(self._h, i, C.byref(name), C.byref(ms), C.byref(steps)) == 0:
            out.append({"kernel": name.value.decode(), "ms_per_step": ms.value / max(steps.value, 1)})
            i += 1
        return out

    @_lib.on_device
    def debug_option(self, option: int, value: int):
        _lib.check(self.lib.iqlb_agents_debug_option(self._h, option, value), "iqlb_agents_debug_option")

    @_lib.on_device
    def debug_tensor(self, which: int) -> torch.Tensor:
        nbytes = C.c_int64()
        _lib.check(self.lib.iqlb_agents_debug_copy(self._h, which, None, C.byref(nbytes), None), "debug")
        out = torch.empty(nbytes.value // 4, dtype=torch.float32, device=self.device)
        _lib.check(self.lib.iqlb_agents_debug_copy(self._h, which, _lib.ptr(out), C.byref(nbytes),
                                                   _lib.stream_ptr(self.device)), "debug")
        return out

    # ---- the reference's checkpoint format, agent by agent (iql.py:664-688, 1519-1521, 1557-1561) ------
    def checkpoint_keys(self, part: str) -> List[Tuple[str, str]]:
        """(packed key, state_dict name) of `part` in {"qf", "vf", "actor"}, in `nn.Module.parameters()` order -
        the order torch's optimizer state is indexed by.  The Gaussian policy registers `log_std` before its
        MLP (iql.py:464-465), so `log_std` is parameter 0 of the actor optimizer."""
        nets = {"qf": (("q1", "q1.net"), ("q2", "q2.net")), "vf": (("v", "v.net"),), "actor": (("actor", "net.net"),)}[part]
        out = [("log_std", "log_std")] if part == "actor" and not self.deterministic else []
        for net, prefix in nets:
            idx = self._actor_lin_idx if net == "actor" else (0, 2, 4)
            for i in range(3):
                out += [(f"{net}.w{i}", f"{prefix}.{idx[i]}.weight"), (f"{net}.b{i}", f"{prefix}.{idx[i]}.bias")]
        return out

    def _optimizer_state_dict(self, agent: int, part: str, lr: float, step: int, fused: Optional[bool]) -> Dict[str, Any]:
        keys = self.checkpoint_keys(part)
        # param_groups as the installed torch writes them for Adam(params, lr=lr[, fused=True]) (iql.py:1481-1491)
        dummy = [nn.Parameter(torch.zeros(1)) for _ in keys]
        sd = torch.optim.Adam(dummy, lr=lr).state_dict()
        sd["param_groups"][0]["fused"] = fused
        state = {}
        if step > 0:
            for i, (key, _) in enumerate(keys):
                state[i] = {"step": torch.tensor(float(step)),
                            "exp_avg": self.view(self.exp_avg, agent, key).detach().clone(),
                            "exp_avg_sq": self.view(self.exp_avg_sq, agent, key).detach().clone()}
        sd["state"] = state
        return sd

    def agent_state_dict(self, agent: int, total_it: Optional[int] = None, fused: Optional[bool] = True) -> Dict[str, Any]:
        """What `ImplicitQLearning.state_dict()` of the reference returns for agent `agent` (iql.py:664-674): the
        three module state_dicts, the three `torch.optim.Adam` state_dicts (state indexed in parameter order), the
        `CosineAnnealingLR` state_dict and `total_it`.  The reference's `load_state_dict` accepts it unchanged."""
        self.sync()
        hp = dict(zip(self.HP_KEYS, self.hp_host[agent]))
        it = int(self.it[agent].item())
        lr_a = float(self.actor_lr[agent].item())
        mod = lambda part: {name: self.view(self.params, agent, key).detach().clone()
                            for key, name in self.checkpoint_keys(part)}
        a_opt = self._optimizer_state_dict(agent, "actor", lr_a, it, fused)
        a_opt["param_groups"][0]["initial_lr"] = hp["actor_lr"]  # added by the scheduler (lr_scheduler.py, _initial_step)
        return {
            "qf": mod("qf"),
            "q_optimizer": self._optimizer_state_dict(agent, "qf", hp["qf_lr"], it, fused),
            "vf": mod("vf"),
            "v_optimizer": self._optimizer_state_dict(agent, "vf", hp["vf_lr"], it, fused),
            "actor": mod("actor"),
            "actor_optimizer": a_opt,
            "actor_lr_schedule": {"T_max": int(hp["max_steps"]), "eta_min": 0.0, "base_lrs": [hp["actor_lr"]],
                                  "last_epoch": it, "_step_count": it + 1, "_is_initial": False,
                                  "_get_lr_called_within_step": False, "_last_lr": [lr_a]},
            "total_it": it if total_it is None else int(total_it),
        }

    def load_agent_state_dict(self, agent: int, state_dict: Dict[str, Any], sync: bool = True) -> int:
        """Resume agent `agent` from a checkpoint of the reference (or of `agent_state_dict`): parameters, Adam
        moments and step, the scheduled actor lr; the target becomes a copy of qf, as iql.py:679 does on load.
        Returns the checkpoint's `total_it`."""
        self.sync()
        parts = (("qf", "q_optimizer"), ("vf", "v_optimizer"), ("actor", "actor_optimizer"))
        mods = {}
        for part, _ in parts:  # validate everything before the first write
            mods[part] = mod = {_strip_compile_prefix_key(k): v for k, v in state_dict[part].items()}
            if part == "actor":
                # nn.Dropout modules have no parameters but shift the Sequential indices: renumber the checkpoint's
                # three Linear layers (sorted by index) onto this engine's naming
                found = sorted({int(k.split(".")[2]) for k in mod if k.startswith("net.net.") and k.endswith(".weight")})
                if len(found) == 3 and tuple(found) != tuple(self._actor_lin_idx):
                    ren = {f"net.net.{a}.": f"net.net.{b}." for a, b in zip(found, self._actor_lin_idx)}
                    mods[part] = mod = {next((ren[p] + k[len(p):] for p in ren if k.startswith(p)), k): v
                                        for k, v in mod.items()}
            missing = [name for _, name in self.checkpoint_keys(part) if name not in mod]
            if missing:
                raise KeyError(f"checkpoint['{part}'] lacks {missing}")
            for key, name in self.checkpoint_keys(part):
                want = tuple(self.view(self.params, agent, key).shape)
                if tuple(mod[name].shape) != want:
                    raise ValueError(f"{part}.{name}: checkpoint shape {tuple(mod[name].shape)} != {want}")
        with torch.no_grad():
            step = 0
            for part, opt in parts:
                mod = mods[part]
                for i, (key, name) in enumerate(self.checkpoint_keys(part)):
                    self.view(self.params, agent, key).copy_(mod[name].to(self.device, torch.float32))
                    st = state_dict[opt]["state"].get(i)
                    m, v = self.view(self.exp_avg, agent, key), self.view(self.exp_avg_sq, agent, key)
                    if st is None:
                        m.zero_(); v.zero_()
                    else:
                        m.copy_(st["exp_avg"].to(self.device, torch.float32))
                        v.copy_(st["exp_avg_sq"].to(self.device, torch.float32))
                        step = int(float(st["step"]))
            self.target[agent, :self.PQ] = self.params[agent, :self.PQ]
            self.it[agent] = step
            self.actor_lr[agent] = float(state_dict["actor_optimizer"]["param_groups"][0]["lr"])
        if sync:
            self._sync_shadow()
        return int(state_dict["total_it"])

    def save_checkpoints(self, paths: Sequence[str], total_it: Optional[int] = None):
        """One reference-format file per agent (`torch.save(trainer.state_dict(), checkpoint_{t}.pt)`, iql.py:1557-1561);
        `paths[i]` is agent i's file."""
        if len(paths) != self.N:
            raise ValueError(f"{self.N} agents need {self.N} paths")
        for a, p in enumerate(paths):
            sd = self.agent_state_dict(a, total_it)
            torch.save(_to_cpu(sd), p)

    def load_checkpoints(self, paths: Sequence[str]) -> List[int]:
        if len(paths) != self.N:
            raise ValueError(f"{self.N} agents need {self.N} paths")
        its = [self.load_agent_state_dict(a, torch.load(p, map_location="cpu", weights_only=False), sync=False)
               for a, p in enumerate(paths)]
        self._sync_shadow()
        return its

    # ---- export in the oracle's naming (tests) -----------------------------------------------------
    def export(self, agent: int) -> Dict[str, np.ndarray]:
        self.sync()
        d = {}
        for key in self._layout:
            d[key] = self.view(self.params, agent, key).detach().cpu().numpy().copy()
            if key.startswith("q1.") or key.startswith("q2."):
                d["qt" + key[1:]] = self.view(self.target, agent, key).detach().cpu().numpy().copy()
        return d


class _ActorLRSchedule:
    """`ImplicitQLearning.actor_lr_schedule` (iql.py:571: `CosineAnnealingLR(actor_optimizer, max_steps)`).  The
    schedule itself advances inside the step kernel (recursive form, torch lr_scheduler.py:1449-1455, one `.step()`
    per `train()` like iql.py:637); this object exposes the scheduler's read / checkpoint surface over the engine's
    `actor_lr[agent]` and `it[agent]`: `get_last_lr()`, `last_epoch`, `T_max`, `eta_min`, `base_lrs`, `state_dict()`,
    `load_state_dict()`.  The actor optimizer's `param_groups[0]["lr"]` is kept in step with it."""

    def __init__(self, trainer: "ImplicitQLearning", agent: int = 0):
        self._tr, self._agent = trainer, agent
        self.eta_min = 0.0

    @property
    def T_max(self) -> int:
        return self._tr._max_steps

    @property
    def base_lrs(self) -> List[float]:
        return [float(self._tr._hp["actor_lr"])]

    @property
    def last_epoch(self) -> int:
        return int(self._tr._engine.it[self._agent].item())

    def get_last_lr(self) -> List[float]:
        return [float(self._tr._engine.actor_lr[self._agent].item())]

    def step(self):
        raise RuntimeError("the cosine schedule is advanced by train() inside the fused step (iql.py:637); "
                           "stepping it separately would advance it twice")

    def state_dict(self) -> Dict[str, Any]:
        it = self.last_epoch
        return {"T_max": self.T_max, "eta_min": 0.0, "base_lrs": self.base_lrs, "last_epoch": it,
                "_step_count": it + 1, "_is_initial": False, "_get_lr_called_within_step": False,
                "_last_lr": self.get_last_lr()}

    def load_state_dict(self, sd: Dict[str, Any]):
        eng = self._tr._engine
        eng.sync()
        eng.actor_lr[self._agent] = float(sd["_last_lr"][0])
        self._tr.actor_optimizer.param_groups[0]["lr"] = float(sd["_last_lr"][0])


class ImplicitQLearning:
    """Drop-in for the reference trainer (iql.py:546-688), one agent on the batched engine."""

    def __init__(self, max_action: float, actor: nn.Module, actor_optimizer: torch.optim.Optimizer,
                 q_network: nn.Module, q_optimizer: torch.optim.Optimizer, v_network: nn.Module,
                 v_optimizer: torch.optim.Optimizer, iql_tau: float = 0.7, beta: float = 3.0,
                 max_steps: int = 1000000, discount: float = 0.99, tau: float = 0.005, device: str = "cuda",
                 batch_size: int = 256, precision: str = "fp32"):
        self.max_action = max_action
        self.qf, self.vf, self.actor = q_network, v_network, actor
        self.v_optimizer, self.q_optimizer, self.actor_optimizer = v_optimizer, q_optimizer, actor_optimizer
        self.iql_tau, self.beta, self.discount, self.tau = iql_tau, beta, discount, tau
        self.device = device
        q_lins = _mlp_linears(q_network.q1)
        _check_two_hidden(q_lins, "q_network")
        a_lins = _mlp_linears(actor.net)
        state_dim, hidden = a_lins[0].in_features, a_lins[0].out_features
        action_dim = a_lins[-1].out_features
        self._max_steps = int(max_steps)
        hp = dict(discount=discount, tau=tau, beta=beta, iql_tau=iql_tau, vf_lr=v_optimizer.param_groups[0]["lr"],
                  qf_lr=q_optimizer.param_groups[0]["lr"], actor_lr=actor_optimizer.param_groups[0]["lr"],
                  max_steps=max_steps)
        self._precision = precision
        self._dims = (state_dim, action_dim, hidden)
        self._hp = hp
        self._engine: Optional[BatchedIQL] = None
        self._batch_size = None
        self.q_target = copy.deepcopy(self.qf).requires_grad_(False).to(device)
        self.actor_lr_schedule = _ActorLRSchedule(self)  # iql.py:571
        self.total_it = 0
        self._build(batch_size)

    def _build(self, batch_size: int):
        s, a, h = self._dims
        det = not hasattr(self.actor, "log_std")
        drop = max([m.p for m in self.actor.modules() if isinstance(m, nn.Dropout)] or [0.0])
        precision = self._precision
        eng = BatchedIQL(s, a, 1, hidden_dim=h, batch_size=batch_size, deterministic=det, device=self.device,
                         precision=precision, hparams=[self._hp], actor_dropout=drop or None)
        eng.adopt(0, self.qf, self.vf, self.actor, bind=True)
        # q_target parameters become views of the packed target buffer
        with torch.no_grad():
            for net, mod in (("q1", self.q_target.q1), ("q2", self.q_target.q2)):
                for i, lin in enumerate(_mlp_linears(mod)):
                    for nm, prm in ((f"{net}.w{i}", lin.weight), (f"{net}.b{i}", lin.bias)):
                        dst = eng.view(eng.target, 0, nm)
                        dst.copy_(prm.detach().to(eng.device))
                        prm.data = dst
        if self._engine is not None:  # batch size changed: carry the optimizer state over
            for name in ("exp_avg", "exp_avg_sq", "it", "actor_lr"):
                getattr(eng, name).copy_(getattr(self._engine, name))
        self._engine = eng
        self._batch_size = batch_size

    @property
    def engine(self) -> BatchedIQL:
        return self._engine

    def train(self, batch: TensorBatch) -> Dict[str, float]:
        self.total_it += 1
        observations, actions, rewards, next_observations, dones = batch
        if observations.shape[0] != self._batch_size:
            self._build(int(observations.shape[0]))
        if actions.shape[1] != self._dims[1]:
            raise RuntimeError("Actions shape missmatch")
        rows = pack_rows(observations, actions, rewards, next_observations, dones, self._engine.R)
        self._engine.train(rows.unsqueeze(0))
        v, q, a = self._engine.losses[0].tolist()  # the reference's three .item() syncs
        return {"value_loss": v, "q_loss": q, "actor_loss": a}

    # ---- checkpoint format of the reference (iql.py:664-688) ------------------------------------------
    def _opt_state(self, opt: torch.optim.Optimizer, keys: Sequence[str], lr: float) -> Dict[str, Any]:
        eng = self._engine
        step = torch.tensor(float(int(eng.it[0].item())))
        state = {}
        for i, key in enumerate(keys):
            state[i] = {"step": step.clone(), "exp_avg": eng.view(eng.exp_avg, 0, key).detach().clone(),
                        "exp_avg_sq": eng.view(eng.exp_avg_sq, 0, key).detach().clone()}
        sd = opt.state_dict()
        if int(eng.it[0].item()) == 0:
            state = {}
        sd["state"] = state
        sd["param_groups"][0]["lr"] = lr
        return sd

    def _keys(self, part: str) -> List[str]:
        """Packed keys of `part` in optimizer (= `parameters()`) order: `log_std` leads the Gaussian actor."""
        return [key for key, _ in self._engine.checkpoint_keys(part)]

    def state_dict(self) -> Dict[str, Any]:
        eng = self._engine
        lr_a = float(eng.actor_lr[0].item())
        it = int(eng.it[0].item())
        return {
            "qf": self.qf.state_dict(),
            "q_optimizer": self._opt_state(self.q_optimizer, self._keys("qf"), self._hp["qf_lr"]),
            "vf": self.vf.state_dict(),
            "v_optimizer": self._opt_state(self.v_optimizer, self._keys("vf"), self._hp["vf_lr"]),
            "actor": self.actor.state_dict(),
            "actor_optimizer": self._opt_state(self.actor_optimizer, self._keys("actor"), lr_a),
            "actor_lr_schedule": self.actor_lr_schedule.state_dict(),
            "total_it": self.total_it,
        }

    def load_state_dict(self, state_dict: Dict[str, Any]):
        """iql.py:676-688 (the target is re-created from qf on load, :679).  The module parameters are views of the
        packed storage, so loading agent slot 0 updates `self.qf / vf / actor / q_target` in place."""
        self.total_it = self._engine.load_agent_state_dict(0, state_dict)


# dataset preparation (iql.py:132-139, 344-401) runs on the GPU; same names and signatures as the reference
from .dataset_prep import (  # noqa: E402,F401
    compute_mean_std, modify_reward, normalize_states, return_reward_range)

# the relabelling half of the reference module lives in relabel.py; re-exported for drop-in imports
from .relabel import (  # noqa: E402,F401
    RewardMLP, RewardPT, _build_mlp_reward_model, _discover_mr_snapshots, _mr_activations, _strip_compile_prefix,
    cvar_stability_check, empirical_cvar, ensemble_predict, load_mlp_reward_model, load_pt_reward_model,
    qlearning_dataset_bnn, qlearning_dataset_mr, qlearning_dataset_mr_ensemble, qlearning_dataset_pt)
