"""Markovian reward MLP.  Mirrors `optbnn/bnn/nets/mlp.py:18-111` (the PyTorch twin that
`algorithms/offline/iql.py:32` imports as RewardMLP) and the call surface of the JAX `Q_MLP`
(`reward_models/q_mlp.py:100-106`: `__call__(observations, actions)`).

The module is a parameter container with the reference's state_dict keys (`layers.0.W`,
`layers.linear_i.W`, `output_layer.W`, W stored (n_in, n_out), N(0,1) init, scaled-variance forward
`X @ (W / sqrt(n_in)) + b`, layers/linear.py:27-47).  `forward`/`predict` run `iqlb_relabel_mlp`.
"""
from typing import List, Sequence

import torch
import torch.nn as nn


class Linear(nn.Module):
    def __init__(self, n_in: int, n_out: int, scaled_variance: bool = True):
        super().__init__()
        if not scaled_variance:
            raise NotImplementedError("only scaled_variance=True (the reference default) is supported")
        self.n_in, self.n_out, self.scaled_variance = n_in, n_out, scaled_variance
        self.W = nn.Parameter(torch.zeros(n_in, n_out))
        self.b = nn.Parameter(torch.zeros(n_out))
        self.reset_parameters()

    def reset_parameters(self):
        nn.init.normal_(self.W, 0, 1.0)
        nn.init.constant_(self.b, 0)


class MLP(nn.Module):
    ACTIVATIONS = ("cos", "tanh", "relu", "softplus", "rbf", "linear", "sin", "leaky_relu", "swish")

    def __init__(self, input_dim: int, output_dim: int, hidden_dims: Sequence[int], activation_fn: str,
                 scaled_variance: bool = True, norm_layer=None, task: str = "regression"):
        super().__init__()
        if activation_fn not in self.ACTIVATIONS:
            raise ValueError(f"unknown activation {activation_fn!r}")
        if norm_layer is not None or task != "regression":
            raise NotImplementedError("relabel path: norm_layer=None, task='regression' only")
        self.input_dim, self.output_dim, self.hidden_dims = input_dim, output_dim, list(hidden_dims)
        self.activation_name = activation_fn
        self.layers = nn.ModuleList([Linear(input_dim, hidden_dims[0], scaled_variance)])
        for i in range(1, len(hidden_dims)):
            self.layers.add_module(f"linear_{i}", Linear(hidden_dims[i - 1], hidden_dims[i], scaled_variance))
        self.output_layer = Linear(hidden_dims[-1], output_dim, scaled_variance)

    def forward(self, X: torch.Tensor, log_softmax: bool = False) -> torch.Tensor:
        from ..relabel import ensemble_predict
        X = X.reshape(-1, self.input_dim)
        out = ensemble_predict([member_weights(self)], X, "mean", activation=self.activation_name, device=X.device)
        return out.unsqueeze(-1)

    def predict(self, X: torch.Tensor) -> torch.Tensor:
        self.eval()
        return self.forward(X)


class Q_MLP(MLP):
    """Call surface of the JAX model: `model(observations, actions)` (reward_models/q_mlp.py:100-106)."""

    def __call__(self, observations, actions=None):
        if actions is None:
            return super().__call__(observations)
        return super().__call__(torch.cat([observations, actions], dim=-1))


def member_weights(model) -> List[torch.Tensor]:
    """[W0, b0, ..., Wout, bout] of a reward MLP (ours or the reference's optbnn MLP)."""
    # the kernels evaluate act(X @ (W / sqrt(n_in)) + b) with no normalisation layer (layers/linear.py:43-47,
    # mlp.py:73-92 with norm_layer=None): a reference MLP built any other way would be relabelled with wrong rewards
    if getattr(model, "norm_layer", None) is not None:
        raise NotImplementedError("reward MLPs with a norm_layer are not supported by the relabel kernels")
    out = []
    for lin in list(model.layers) + [model.output_layer]:
        if getattr(lin, "scaled_variance", True) is False:
            raise NotImplementedError("reward MLPs with scaled_variance=False are not supported by the relabel kernels")
        out += [lin.W.detach(), lin.b.detach()]
    if out[-2].shape[1] != 1:
        raise ValueError("reward MLP must have output_dim == 1")
    return out
