"""Import the UNMODIFIED reference (`/root/reference`) in the authoring container.

TEST INFRASTRUCTURE ONLY.  Used by `oracle/gen_golden.py` (fixture generation) and by
CPU tests that pin the oracle restatement against the real reference when
`/root/reference` exists.  Nothing on the product path, and nothing that runs on the
GPU box, may import this module (the reference does not travel).

The reference imports third-party modules that are absent here (d4rl, gym, pyrallis,
h5py, matplotlib, prettytable, seaborn).  None of them is touched by the hot path, so
they are replaced by empty stand-ins in `sys.modules`.
"""
import importlib
import importlib.util
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("IQLPREF_REFERENCE_ROOT", "/root/reference")


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
