"""ctypes binding of the C ABI in ``include/lgasr.h`` (``csrc/liblgasr.so``).

The library is CUDA-only.  If it has not been built, or no CUDA device is present when a compute
entry point is called, this module raises: there is no CPU fallback and no eager-PyTorch path.
"""

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "liblgasr.so")

MAX_TERMS = 8
MAX_PREBEAM = 128


class Config(C.Structure):
    _fields_ = [
        ("n_utts", C.c_int32),
        ("max_frames", C.c_int32),
        ("vocab", C.c_int32),
        ("beam", C.c_int32),
        ("pre_beam", C.c_int32),
        ("n_terms", C.c_int32),
        ("pre_beam_key", C.c_int32),
        ("use_ctc", C.c_int32),
        ("ctc_weight", C.c_float),
        ("blank", C.c_int32),
        ("sos", C.c_int32),
        ("eos", C.c_int32),
        ("max_len_cap", C.c_int32),
        ("primer_len", C.c_int32),
        ("end_detect", C.c_int32),
    ]


_P = C.c_void_p
_I = C.c_int
_I64 = C.c_int64
_PI32 = C.POINTER(C.c_int32)
_PF = C.POINTER(C.c_float)
_PI64 = C.POINTER(C.c_int64)

# name -> (restype, argtypes); every symbol include/lgasr.h declares
PROTOTYPES = {
    "lgasr_version": (_I, []),
    "lgasr_last_error": (C.c_char_p, []),
    "lgasr_launch_count": (_I64, []),
    "lgasr_create": (_I, [C.POINTER(Config), C.POINTER(_P)]),
    "lgasr_destroy": (_I, [_P]),
    "lgasr_set_posteriors": (_I, [_P, _P, _PI32, _P]),
    "lgasr_set_posteriors_host": (_I, [_P, _P, _PI32, _P]),
    "lgasr_begin": (_I, [_P, _PI32, _PI32, _PI64, _P]),
    "lgasr_refill": (_I, [_P, _I, _P, _I, _I, _I, _P]),
    "lgasr_step": (_I, [_P, C.POINTER(_P), _PF, _PF, _I64, _P]),
    "lgasr_view": (_I, [_P, _I, C.POINTER(_P), C.POINTER(_I64)]),
    "lgasr_poll": (_I, [_P, _PI32, _PI32, _PI32, _PI32, _PI32, _P]),
    "lgasr_gather_ended": (_I, [_P, _I, _PI32, _PI32, _PI64, _PI32, _PF, _PF, _P]),
    "lgasr_gather_best": (_I, [_P, _I, _I, _PI64, _PI32, _PF, _P]),
    "lgasr_launch_info": (_I, [_P, _PI32, _PI32, _PI32, _PI32]),
    "lgasr_set_profiling": (_I, [_P, _I]),
    "lgasr_get_profile": (_I, [_P, _PF, _PF, _PI32]),
    "lgasr_debug_trace": (_I, [_P, _PI64, _I64]),
    "lgasr_decode_replay": (_I, [_P, _I, C.POINTER(_P), _PF, _PF, _I64, _P]),
    "lgasr_decode_replay_host": (_I, [_P, _P, _PI32, _PI32, _PI32, _PI64, _I, C.POINTER(_P), _PF, _PF, _I64, _P]),
    "lgasr_transpose_posteriors": (_I, [_P, _I, _I, _I, _I, _P, _I, _P]),
    "lgasr_ctc_initial_state": (_I, [_P, _I, _I, _I, _P, _P]),
    "lgasr_ctc_prefix_score": (_I, [_P, _I, _I, _I, _I, _I, _I, _P, _P, _P, _I, _P, _I, _P, _P, _P]),
    "lgasr_ctc_prefix_score_win": (_I, [_P, _I, _I, _I, _I, _I, _I, _P, _P, _P, _I, _P, _I, _P, _P, _I, _I, _P]),
    "lgasr_ctc_select_state_win": (_I, [_P, _I, _I, _I, _I, _P, _P, _P, _P, _I, _P, _I, _I, _P]),
    "lgasr_ctc_extend_state": (_I, [_P, _I, _I, _I, _I, _I, _I, _P, _P, _P]),
    "lgasr_ctc_select_state": (_I, [_P, _I, _I, _I, _I, _P, _P, _P, _P, _I, _P, _P]),
    "lgasr_reorder_rows": (_I, [_P, _P, _P, _I64, _I64, _I64, _I64, _I64, _I64, _P]),
    "lgasr_log_softmax_rows": (_I, [_P, _I, _I64, _P, _I64, _I, _I, _P]),
    "lgasr_row_logsumexp": (_I, [_P, _I, _I64, _P, _I, _I, _P]),
    "lgasr_row_logsumexp_noalloc": (_I, [_P, _I, _I64, _P, _I, _I, _P]),
    "lgasr_step_ex": (_I, [_P, C.POINTER(_P), _PI32, C.POINTER(_P), _PF, _PF, _I64, _P]),
    "lgasr_score_filter": (_I, [_P, _I64, _I, _I, _I, _I, _I, _I, _I, _I, _P, _I64, _P]),
}

_lib = None


class LgasrError(RuntimeError):
    pass


def load():
    """Load ``liblgasr.so`` (once). Raises if it has not been built: no fallback exists."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise LgasrError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C llm_guided_asr_b200/csrc`. This package has no CPU / eager fallback."
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export it
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        msg = load().lgasr_last_error()
        raise LgasrError(f"lgasr error {rc}: {msg.decode() if msg else '?'}")


def launch_count() -> int:
    return int(load().lgasr_launch_count())


class DeviceView:
    """Zero-copy view of a device buffer owned by a handle, via ``__cuda_array_interface__``."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {
            "shape": tuple(shape),
            "typestr": typestr,
            "data": (int(ptr), False),
            "version": 2,
            "strides": None,
        }


def i32_array(values):
    arr = (C.c_int32 * len(values))(*[int(v) for v in values])
    return arr


def i64_array(values):
    return (C.c_int64 * len(values))(*[int(v) for v in values])


def f32_array(values):
    return (C.c_float * len(values))(*[float(v) for v in values])


def ptr_array(values):
    return (_P * len(values))(*[(_P(int(v)) if v else _P(0)) for v in values])
