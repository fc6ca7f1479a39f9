"""Import the UNMODIFIED reference (`/root/reference`) in the authoring container.

TEST / BENCH INFRASTRUCTURE ONLY.  Used by `oracle/gen_golden.py` (fixture generation), by tests
that pin the oracle restatement and the CUDA path against the real reference, and by
`bench.py --impl reference` / `reference_cuda`.  Nothing on the product path may import this
module.  On the GPU box `/root/reference` does not exist: the copy of the hot-path files under
`baseline/_ref/` (oracle/make_ref.py; git-ignored, shipped) is used when present, and every
caller falls back to the committed fixtures / the oracle port when it is not.

The reference imports third-party modules that are absent here (d4rl, gym, pyrallis,
h5py, matplotlib, prettytable, seaborn).  None of them is touched by the hot path, so
they are replaced by empty stand-ins in `sys.modules`.
"""
import importlib
import importlib.util
import os
import sys
import types

_REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _find_root() -> str:
    """`/root/reference` in the authoring container; on the GPU box the git-ignored copy of the hot-path files that
    `oracle/make_ref.py` placed under `baseline/_ref/` (it ships with the snapshot, the reference itself does not)."""
    for root in (os.environ.get("IQLPREF_REFERENCE_ROOT"), "/root/reference", os.path.join(_REPO, "baseline", "_ref")):
        if root and os.path.isfile(os.path.join(root, "algorithms/offline/iql.py")):
            return root
    return "/root/reference"


REFERENCE_ROOT = _find_root()


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "algorithms/offline/iql.py"))


def _stub(name: str, **attrs):
    if name in sys.modules:
        return sys.modules[name]
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def _install_stubs():
    def _identity_wrap(*a, **k):
        def deco(fn):
            return fn
        return deco

    gym = _stub("gym", Env=object)
    gym.wrappers = _stub("gym.wrappers")
    gym.vector = _stub("gym.vector")
    _stub("d4rl")
    _stub("pyrallis", wrap=_identity_wrap, dump=lambda *a, **k: None)
    _stub("h5py")
    mpl = _stub("matplotlib")
    mpl.pyplot = _stub("matplotlib.pyplot")
    _stub("prettytable", PrettyTable=object)
    _stub("seaborn")
    if "wandb" not in sys.modules:
        try:
            importlib.import_module("wandb")
        except Exception:
            _stub("wandb")


_CACHE = {}


def load_reference_iql():
    """Return the reference module `algorithms/offline/iql.py` (unmodified)."""
    if "iql" in _CACHE:
        return _CACHE["iql"]
    if not reference_available():
        raise FileNotFoundError(f"reference not found under {REFERENCE_ROOT}")
    _install_stubs()
    optbnn_root = os.path.join(REFERENCE_ROOT, "gp_reward-priors")
    if optbnn_root not in sys.path:
        sys.path.insert(0, optbnn_root)
    path = os.path.join(REFERENCE_ROOT, "algorithms/offline/iql.py")
    spec = importlib.util.spec_from_file_location("_reference_iql", path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules["_reference_iql"] = mod
    spec.loader.exec_module(mod)
    _CACHE["iql"] = mod
    return mod
