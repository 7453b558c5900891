"""ctypes binding of libb200rl.so (C-ABI declared in include/b200rl.h).

The library is built in-tree by `make -C handyrl_b200/csrc` (or
`__graft_entry__.build()`).  Loading fails loudly when it is missing: the
product has no CPU or eager-PyTorch fallback.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# B200RL_LIB selects another build of the same library (kernel A/B experiments)
LIB_PATH = os.environ.get("B200RL_LIB") or os.path.join(_HERE, "libb200rl.so")

_lib = None

c_i64 = ctypes.c_int64
c_int = ctypes.c_int
c_ptr = ctypes.c_void_p
c_dbl = ctypes.c_double
c_u64 = ctypes.c_uint64


class B200RLError(RuntimeError):
    pass


class TargetHead(ctypes.Structure):
    """b200rl_target_head (include/b200rl.h)."""
    _fields_ = ([("values", c_ptr), ("v_sb", c_i64), ("v_st", c_i64), ("v_sp", c_i64),
                 ("boot", c_ptr), ("boot_sb", c_i64), ("boot_sp", c_i64),
                 ("rewards", c_ptr), ("r_sb", c_i64), ("r_st", c_i64), ("r_sp", c_i64),
                 ("rhos", c_ptr), ("rho_sb", c_i64), ("rho_st", c_i64), ("rho_sp", c_i64),
                 ("cs", c_ptr), ("c_sb", c_i64), ("c_st", c_i64), ("c_sp", c_i64),
                 ("out_targets", c_ptr), ("out_adv", c_ptr), ("gamma", c_dbl)])


def _view3():
    return [c_ptr, c_i64, c_i64, c_i64]


_SCAN_TAIL = (
    _view3()                      # values
    + [c_ptr, c_i64, c_i64]       # boot
    + _view3()                    # rewards
    + _view3()                    # rhos
    + _view3()                    # cs
    + [c_ptr, c_ptr, c_i64, c_i64, c_i64, c_dbl, c_dbl, c_ptr]
)

# name -> argtypes; restype is int except where noted.  Must list every symbol
# include/b200rl.h declares (tests/test_abi.py checks this against the header).
SIGNATURES = {
    "b200rl_version": [],
    "b200rl_target_scan": [c_int] + _SCAN_TAIL,
    "b200rl_target_scan_fused": [c_int, c_int] + _SCAN_TAIL,
    "b200rl_target_scan_heads": [c_int, c_int, c_ptr, c_int, c_i64, c_i64, c_i64, c_dbl, c_ptr],
    "b200rl_target_mc": _view3() + _view3() + [c_ptr, c_i64, c_i64, c_i64, c_ptr],
    "b200rl_env_state_bytes": [c_int],
    "b200rl_env_dims": [c_int, c_ptr],
    "b200rl_env_reset": [c_int, c_ptr, c_i64, c_ptr, c_ptr],
    "b200rl_env_step": [c_int, c_ptr, c_i64, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr],
    "b200rl_env_legal_list": [c_int, c_ptr, c_i64, c_ptr, c_ptr, c_ptr],
    "b200rl_env_observe": [c_int, c_ptr, c_i64, c_int, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr],
    "b200rl_env_outcome": [c_int, c_ptr, c_i64, c_ptr, c_ptr],
    "b200rl_env_info": [c_int, c_ptr, c_i64, c_ptr, c_ptr],
    "b200rl_env_step_observe": [c_int, c_ptr, c_i64, c_ptr, c_ptr, c_int] + [c_ptr] * 8,
    "b200rl_env_rollout": [c_int, c_ptr, c_i64, c_i64, c_u64, c_u64, c_u64, c_ptr, c_ptr, c_ptr, c_ptr, c_i64,
                           c_ptr, c_i64, c_ptr, c_i64, c_ptr, c_ptr, c_i64, c_ptr],
    "b200rl_env_rollout_scratch_bytes": [c_i64],
    "b200rl_env_rollout_static": [c_int, c_ptr, c_i64, c_i64, c_u64, c_u64, c_u64, c_ptr, c_ptr, c_ptr, c_ptr, c_i64,
                                  c_ptr, c_i64, c_ptr, c_i64, c_ptr, c_ptr],
    "b200rl_set_tuning": [ctypes.c_char_p, c_i64, c_int],
    "b200rl_env_step_record": [c_int, c_ptr, c_i64, c_ptr, c_ptr, c_ptr, c_int, c_ptr, c_i64, c_u64, c_ptr, c_i64,
                               c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr],
    "b200rl_env_logit_slots": [c_int],
    "b200rl_env_sample_record": [c_int, c_ptr, c_i64, c_ptr, c_ptr, c_u64, c_u64, c_ptr, c_i64, c_u64, c_ptr, c_i64,
                                 c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr],
    "b200rl_make_batch": [c_int, c_ptr, c_i64, c_i64, c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_int, c_int, c_ptr,
                          c_i64] + [c_ptr] * 14,
    "b200rl_select_windows": [c_ptr, c_i64, c_ptr, c_i64, c_i64, c_u64, c_u64, c_u64, c_i64, c_ptr, c_i64, c_ptr],
    "b200rl_select_windows_host": [c_ptr, c_i64, c_u64, c_i64, c_i64, c_u64, c_u64, c_u64, c_i64, c_ptr, c_i64, c_int],
    "b200rl_policy_terms": [c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_dbl, c_dbl, c_ptr, c_ptr, c_ptr, c_ptr],
    "b200rl_policy_terms_backward": [c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_ptr],
    "b200rl_value_blend": [c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_i64, c_int, c_ptr, c_ptr],
    "b200rl_pad_stack": [c_ptr, c_ptr, c_ptr, ctypes.c_uint32, c_ptr, c_i64, c_i64, c_i64, c_ptr],
}

# functions that return a value instead of a status
_RESTYPE = {"b200rl_env_state_bytes": c_i64, "b200rl_env_logit_slots": c_i64,
            "b200rl_env_rollout_scratch_bytes": c_i64}


def lib():
    """Load (once) and return the ctypes handle of libb200rl.so."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B200RLError(
            "libb200rl.so not found at %s — build it with `make -C handyrl_b200/csrc` "
            "(there is no CPU fallback)" % LIB_PATH
        )
    h = ctypes.CDLL(LIB_PATH)
    h.b200rl_last_error.restype = ctypes.c_char_p
    h.b200rl_last_error.argtypes = []
    for name, argtypes in SIGNATURES.items():
        fn = getattr(h, name)
        fn.argtypes = argtypes
        fn.restype = _RESTYPE.get(name, c_int)
    _lib = h
    return h


def check(status, what=""):
    if status != 0:
        msg = lib().b200rl_last_error()
        raise B200RLError("%s failed (%d): %s" % (what, status, msg.decode() if msg else "?"))


def call(name, *args):
    check(getattr(lib(), name)(*args), name)


def set_tuning(name, value=None):
    """Set (or, with value=None, clear) a tuning knob of the library (include/b200rl.h: b200rl_set_tuning)."""
    call("b200rl_set_tuning", name.encode(), 0 if value is None else int(value), 0 if value is None else 1)


def stream_ptr():
    import torch

    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def require_cuda(t, name):
    if not t.is_cuda:
        raise B200RLError(
            "%s must be a CUDA tensor: handyrl_b200 runs only on the GPU (no CPU fallback)" % name
        )
