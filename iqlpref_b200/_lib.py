"""ctypes binding of libiqlb.so (C ABI in include/iqlb.h).

There is no CPU fallback: if the shared library is missing or a call fails, this raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("IQLB_LIB") or os.path.join(_HERE, "csrc", "libiqlb.so")

HP_STRIDE = 8

ACTIVATIONS = {"relu": 0, "tanh": 1, "cos": 2, "sin": 3, "linear": 4, "rbf": 5, "swish": 6, "softplus": 7,
               "leaky_relu": 8}


class IqlbError(RuntimeError):
    pass


class AgentsDesc(C.Structure):
    _fields_ = [("n_agents", C.c_int32), ("state_dim", C.c_int32), ("action_dim", C.c_int32),
                ("hidden_dim", C.c_int32), ("batch_size", C.c_int32), ("deterministic", C.c_int32),
                ("row_stride", C.c_int32), ("precision", C.c_int32)]


class PTDesc(C.Structure):
    _fields_ = [("state_dim", C.c_int32), ("action_dim", C.c_int32), ("embd_dim", C.c_int32),
                ("num_heads", C.c_int32), ("intermediate_dim", C.c_int32), ("num_layers", C.c_int32),
                ("pref_dim", C.c_int32), ("n_timesteps", C.c_int32), ("query_len", C.c_int32), ("eps", C.c_float)]


_P = C.c_void_p
_SIGNATURES = {
    "iqlb_version": ([], C.c_int),
    "iqlb_last_error": ([], C.c_char_p),
    "iqlb_philox_randint": ([C.c_uint64, C.c_uint64, C.c_int64, C.c_int32, _P, _P], C.c_int),
    "iqlb_replay_sample": ([_P, C.c_int64, C.c_int32, C.c_int32, _P, _P, _P, C.c_int32, C.c_int32, _P, _P, _P], C.c_int),
    "iqlb_replay_gather": ([_P, C.c_int64, C.c_int32, C.c_int32, _P, _P, _P, C.c_int32, C.c_int32, _P, _P], C.c_int),
    "iqlb_param_count": ([C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_int64), C.POINTER(C.c_int64)], C.c_int),
    "iqlb_agents_create": ([C.POINTER(AgentsDesc), _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, C.POINTER(_P)], C.c_int),
    "iqlb_agents_destroy": ([_P], C.c_int),
    "iqlb_agents_set_replay": ([_P, _P, C.c_int64, _P, _P, _P], C.c_int),
    "iqlb_agents_step": ([_P, _P, _P], C.c_int),
    "iqlb_agents_step_host": ([_P, _P, _P, _P], C.c_int),
    "iqlb_agents_step_host_pipelined": ([_P, _P, _P, _P], C.c_int),
    "iqlb_agents_refresh": ([_P, _P], C.c_int),
    "iqlb_agents_set_overlap": ([_P, C.c_int32, _P], C.c_int),
    "iqlb_agents_sync": ([_P, _P], C.c_int),
    "iqlb_agents_launches_per_step": ([_P, C.c_int32], C.c_int),
    "iqlb_agents_profile": ([_P, C.c_int32], C.c_int),
    "iqlb_agents_profile_collect": ([_P], C.c_int),
    "iqlb_agents_profile_read": ([_P, C.c_int32, C.POINTER(C.c_char_p), C.POINTER(C.c_double), C.POINTER(C.c_int32)], C.c_int),
    "iqlb_agents_debug_option": ([_P, C.c_int32, C.c_int32], C.c_int),
    "iqlb_agents_set_actor_dropout": ([_P, C.c_float], C.c_int),
    "iqlb_agents_debug_copy": ([_P, C.c_int32, _P, C.POINTER(C.c_int64), _P], C.c_int),
    "iqlb_agents_act": ([_P, _P, C.c_int64, C.c_int32, C.c_float, _P, _P], C.c_int),
    "iqlb_polyak": ([_P, _P, C.c_int64, C.c_float, _P], C.c_int),
    "iqlb_relabel_mlp": ([_P, C.c_int64, C.c_int32, C.c_int32, _P, C.c_int32, C.POINTER(C.c_int32), C.c_int32,
                          C.c_int32, C.c_int32, _P, _P, _P, _P], C.c_int),
    "iqlb_relabel_mlp_p": ([_P, C.c_int64, C.c_int32, C.c_int32, _P, C.c_int32, C.POINTER(C.c_int32), C.c_int32,
                            C.c_int32, C.c_int32, C.c_int32, _P, _P, _P, _P, _P], C.c_int),
    "iqlb_relabel_debug_timeline": ([_P], C.c_int),
    "iqlb_pt_param_count": ([C.POINTER(PTDesc), C.POINTER(C.c_int64)], C.c_int),
    "iqlb_relabel_pt": ([C.POINTER(PTDesc), _P, _P, _P, C.c_int64, _P, _P, C.c_int64, _P, _P], C.c_int),
    "iqlb_pt_forward": ([C.POINTER(PTDesc), _P, _P, _P, _P, _P, C.c_int64, _P, _P, _P], C.c_int),
    "iqlb_selftest_umma": ([_P, C.c_int64, _P, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                            C.c_int32, C.c_int32, _P, _P], C.c_int),
    "iqlb_selftest_gemm": ([C.c_int32, _P, C.c_int32, _P, C.c_int32, _P, _P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                            C.c_int32, C.c_float, C.c_int32, _P], C.c_int),
    "iqlb_selftest_mma_rate": ([C.c_int32, C.c_int32, C.c_int32, _P, _P], C.c_int),
    "iqlb_reward_range": ([_P, _P, C.c_int64, C.c_int32, _P, _P, _P], C.c_int),
    "iqlb_reward_range_static": ([_P, _P, C.c_int64, C.c_int32, _P, _P, _P], C.c_int),
    "iqlb_modify_reward": ([_P, _P, C.c_int64, C.c_int32, C.c_int32, C.c_int32, _P], C.c_int),
    "iqlb_state_mean_std": ([_P, C.c_int64, C.c_int32, C.c_int32, C.c_float, _P, _P, _P], C.c_int),
    "iqlb_normalize_states": ([_P, C.c_int64, C.c_int32, C.c_int32, _P, _P, _P], C.c_int),
    "iqlb_mr_create": ([C.c_int32, C.POINTER(C.c_int32), C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.POINTER(_P)], C.c_int),
    "iqlb_mr_destroy": ([_P], C.c_int),
    "iqlb_mr_set_precision": ([_P, C.c_int32], C.c_int),
    "iqlb_mr_param_count": ([_P, C.POINTER(C.c_int64)], C.c_int),
    "iqlb_mr_train_step": ([_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                            C.c_float, C.c_float, C.c_float, C.c_float, C.c_int64, _P, _P, _P], C.c_int),
    "iqlb_mr_train_step_f": ([_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                              C.c_float, C.c_float, C.c_float, C.c_float, C.c_int64, _P, _P, C.c_int32, C.c_int32, _P, _P, _P,
                              C.c_double, C.c_double, C.c_double, _P, _P, _P, _P], C.c_int),
    "iqlb_mr_eval": ([_P, _P, _P, _P, _P, _P, _P, _P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_int32, _P, _P, _P], C.c_int),
    "iqlb_pt_train_param_count": ([C.POINTER(PTDesc), C.POINTER(C.c_int64)], C.c_int),
    "iqlb_pt_trainer_create": ([C.POINTER(PTDesc), C.c_int32, C.c_int32, C.POINTER(_P)], C.c_int),
    "iqlb_pt_trainer_destroy": ([_P], C.c_int),
    "iqlb_pt_trainer_set_precision": ([_P, C.c_int32], C.c_int),
    "iqlb_pt_trainer_set_dropout": ([_P, C.c_float, C.c_float, C.c_float, C.c_float, C.c_uint64], C.c_int),
    "iqlb_pt_train_step": ([_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, C.c_int32, C.c_float, C.c_float, C.c_float,
                            C.c_float, C.c_int64, C.c_int32, _P, _P, _P], C.c_int),
    "iqlb_pt_eval": ([_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, C.c_int32, _P, _P, _P], C.c_int),
    "iqlb_keep_mask_scan": ([_P, _P, C.c_int64, C.c_int32, C.c_int32, _P, _P], C.c_int),
}

EXPORTS = tuple(_SIGNATURES)
_lib = None


def load():
    """Load libiqlb.so (once).  Raises IqlbError when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise IqlbError(f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                        "(there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, (argtypes, restype) in _SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        fn.argtypes = argtypes
        fn.restype = restype
    if lib.iqlb_version() != 100:
        raise IqlbError(f"libiqlb.so version {lib.iqlb_version()} does not match the Python binding (100)")
    _lib = lib
    return lib


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = load().iqlb_last_error().decode(errors="replace")
        if rc == -1:
            raise ValueError(f"{what}: {msg}")
        raise IqlbError(f"{what}: iqlb error {rc}: {msg}")


def on_device(fn):
    """Method decorator: run with `self.device` as the current CUDA device.  The library allocates its scratch on, and
    launches on, the CURRENT device, while the buffers and the stream it is handed belong to `self.device`; without
    this a caller using device="cuda:1" and no prior torch.cuda.set_device would mix devices."""
    import functools

    @functools.wraps(fn)
    def wrapper(self, *a, **k):
        import torch
        with torch.cuda.device(self.device):
            return fn(self, *a, **k)
    return wrapper


def device_ctx(device):
    """`with device_ctx(dev):` for free functions (same purpose as `on_device`)."""
    import torch
    return torch.cuda.device(device)


def ptr(t):
    """Raw device/host pointer of a torch tensor (None -> NULL)."""
    return None if t is None else C.c_void_p(t.data_ptr())


def stream_ptr(device=None):
    import torch
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)
