"""ctypes binding of libhm_engine.so — the C ABI of include/hm_engine.h.

There is no CPU fallback: if the library is missing or cannot be loaded this module raises.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libhm_engine.so")

NB_INPUT_VALUES = 4736
NB_POLICY_VALUES = 4672
DTYPE_F32, DTYPE_F16, DTYPE_BF16, DTYPE_U8 = 0, 1, 2, 3


class ValueConfig(C.Structure):
    _fields_ = [
        ("enable_wdl_eval", C.c_int),
        ("wdl_value_weight", C.c_float),
        ("moves_left_discount", C.c_float),
        ("draw_contempt", C.c_float),
    ]


class RootPolicy(C.Structure):  # hm_root_policy
    _fields_ = [
        ("temperature", C.c_double),
        ("dirichlet_alpha", C.c_float),
        ("dirichlet_epsilon", C.c_float),
        ("noise_seed", C.c_uint64),
        ("position_hash", C.c_void_p),
        ("captures", C.c_void_p),
        ("uniforms", C.c_void_p),
    ]


ROOT_NOALLOC_MAX_MOVES = 1 << 20  # HM_ROOT_NOALLOC_MAX_MOVES
ROOT_OK, ROOT_RESAMPLED, ROOT_RESAMPLED_FORCED = 0, 1, 2
ROOT_EMPTY_LIST, ROOT_BAD_WEIGHTS, ROOT_NO_LEGAL_JOINT = -1, -2, -3


class KernelTime(C.Structure):
    _fields_ = [("name", C.c_char * 48), ("ms", C.c_float), ("flops", C.c_double), ("bytes", C.c_double)]


# every symbol include/hm_engine.h declares: name -> (restype, argtypes)
_vp, _sz, _i = C.c_void_p, C.c_size_t, C.c_int
_pp = C.POINTER(C.c_void_p)
SYMBOLS = {
    "hm_engine_create": (_vp, [_i, _i, _i]),
    "hm_engine_destroy": (None, [_vp]),
    "hm_engine_load": (_i, [_vp, C.c_char_p]),
    "hm_engine_load_memory": (_i, [_vp, _vp, _sz]),
    "hm_engine_batch_size": (_i, [_vp]),
    "hm_engine_num_workers": (_i, [_vp]),
    "hm_engine_arch": (_i, [_vp]),
    "hm_engine_enqueue_half": (_i, [_vp, _vp, _sz]),
    "hm_engine_sync_half": (_i, [_vp, _sz, _pp, _pp, _pp, _pp, _pp]),
    "hm_engine_run_half": (_i, [_vp, _vp, _sz, _pp, _pp, _pp, _pp, _pp]),
    "hm_engine_run_float": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz]),
    "hm_engine_enqueue_positions": (_i, [_vp, _vp, _sz, _vp, _vp, _sz]),
    "hm_engine_sync_priors": (_i, [_vp, _sz, _pp, _pp, _pp]),
    "hm_root_policy_apply": (_i, [_vp, _vp, _sz, _vp, _vp, _vp, C.POINTER(RootPolicy), _vp, _vp, _vp]),
    "hm_root_policy_apply_noalloc": (_i, [_i, _vp, _sz, _vp, _vp, _sz, _vp, C.POINTER(RootPolicy), _vp, _vp, _vp]),
    "hm_engine_enqueue_positions_root": (_i, [_vp, _vp, _sz, _vp, _vp, C.POINTER(RootPolicy), _sz]),
    "hm_engine_sync_root": (_i, [_vp, _sz, _pp, _pp, _pp, _pp, _pp]),
    "hm_engine_set_value_config": (C.c_int, [_vp, C.POINTER(ValueConfig)]),
    "hm_encode_planes": (_i, [_vp, _vp, _sz, _vp, _i]),
    "hm_policy_priors_half": (_i, [_vp, _vp, _vp, _vp, _sz, _vp, _vp, _vp]),
    "hm_policy_indices": (_i, [_vp, _vp, _vp, _sz, _vp]),
    "hm_shape_values_half": (_i, [_vp, _vp, _vp, _vp, _sz, _vp]),
    "hm_engine_bench_e2e": (_i, [_vp, _vp, _sz, _vp, _vp, _i, _i, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "hm_bench_planes": (_i, [_vp, _vp, _sz, _i, _i, C.POINTER(C.c_float)]),
    "hm_bench_priors": (_i, [_vp, _vp, _sz, _vp, _vp, _i, C.POINTER(C.c_float)]),
    "hm_engine_bench_resident": (_i, [_vp, _sz, _i, C.POINTER(C.c_float), C.POINTER(_i)]),
    "hm_engine_trace_trunk": (_i, [_vp, _sz, _i, _vp, C.POINTER(_i)]),
    "hm_engine_bench_resident_multi": (_i, [_vp, _i, _i, C.POINTER(C.c_float), C.POINTER(_i)]),
    "hm_engine_profile_pass": (_i, [_vp, _sz, C.POINTER(KernelTime), _i, C.POINTER(_i)]),
    "hm_engine_debug_fetch": (_i, [_vp, _sz, C.c_char_p, _vp, _sz, C.POINTER(_sz)]),
    "hm_last_error": (C.c_char_p, []),
    "hm_version": (C.c_char_p, []),
    "hm_alloc_pinned": (_vp, [_sz]),
    "hm_free_pinned": (None, [_vp]),
}

_lib = None


def load() -> C.CDLL:
    """Load libhm_engine.so (building is explicit: ``python -m hivemind_b200.build``)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -m hivemind_b200.build` "
            "(there is no CPU fallback)"
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the ABI is incomplete
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error() -> str:
    return load().hm_last_error().decode(errors="replace")
