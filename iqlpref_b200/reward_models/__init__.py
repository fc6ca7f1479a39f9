"""Reward models of the relabelling pass (mirrors `reward_models/` and the torch twins in
`gp_reward-priors/optbnn/bnn/nets/` of the reference)."""
from .q_mlp import MLP, Q_MLP  # noqa: F401
from .pref_transformer import PT  # noqa: F401
