"""ctypes binding of include/wrs.h.  Loads the in-tree libwrs_b200.so; there is no fallback of any kind:
if the library is missing or a call fails, an exception is raised."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "lib", "libwrs_b200.so")

WRS_OK, WRS_ERR_INVALID, WRS_ERR_CUDA, WRS_ERR_UNSUPPORTED, WRS_ERR_NUMERIC, WRS_ERR_TIMEOUT = range(6)
FILTER_SUPPRESS_BLANK, FILTER_SUPPRESS_TOKENS, FILTER_TIMESTAMPS, FILTER_ALL = 1, 2, 4, 7
ROW_TS_MASKED, ROW_TEXT_LT_EOT_MASKED, ROW_FIRST_STEP, ROW_TS_OVER_TEXT, ROW_NAN = 1, 2, 4, 8, 128


class WrsError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"wrs error {code}: {msg}")
        self.code = code


class DecodeCfg(C.Structure):
    _fields_ = [("n_vocab", C.c_int32), ("eot", C.c_int32), ("sot", C.c_int32), ("no_timestamps", C.c_int32),
                ("ts_begin", C.c_int32), ("no_speech", C.c_int32), ("blank_ids", C.c_int32 * 2),
                ("suppress_blank", C.c_int32), ("timestamps", C.c_int32), ("max_initial_timestamp_index", C.c_int32),
                ("beam_size", C.c_int32), ("patience", C.c_float), ("max_len", C.c_int32)]


class StepDebug(C.Structure):
    _fields_ = [("cand_logprob", C.c_void_p), ("cand_token", C.c_void_p), ("row_flags", C.c_void_p),
                ("fin_count", C.c_void_p), ("fin_source", C.c_void_p), ("fin_score", C.c_void_p)]


class KvTensor(C.Structure):
    """wrs_kv_tensor (include/wrs.h): one cached K or V array, rows contiguous."""
    _fields_ = [("src", C.c_void_p), ("dst", C.c_void_p), ("row_bytes", C.c_int64), ("group_invariant", C.c_int32),
                ("reserved", C.c_int32)]


# name -> (restype, argtypes); every symbol declared in include/wrs.h
_P, _I32, _I64, _F, _D = C.c_void_p, C.c_int32, C.c_int64, C.c_float, C.c_double
SIGNATURES = {
    "wrs_abi_version": (C.c_int, []),
    "wrs_ctx_create": (C.c_int, [C.c_int, C.POINTER(_P)]),
    "wrs_ctx_destroy": (None, [_P]),
    "wrs_last_error": (C.c_char_p, [_P]),
    "wrs_ctx_check": (C.c_int, [_P, _P]),
    "wrs_mel_filter_host": (C.c_int, [_P]),
    "wrs_logmel_set_filter": (C.c_int, [_P, _P]),
    "wrs_logmel_batch": (C.c_int, [_P, _P, _I64, _I64, _I64, _P, _P]),
    "wrs_logmel_batch_noalloc": (C.c_int, [_P, _P, _I64, _I64, _I64, _P, _P]),
    "wrs_logmel_batch_host": (C.c_int, [_P, _P, _I64, _I64, _I64, _P]),
    "wrs_pcm_i16_to_f32": (C.c_int, [_P, _P, _I64, _I32, _P, _P]),
    "wrs_logmel_batch_host_i16": (C.c_int, [_P, _P, _I64, _I64, _I64, _P]),
    "wrs_pad_or_trim": (C.c_int, [_P, _P, _I64, _I64, _I64, _P, _I64, _P]),
    "wrs_ctx_set_option": (C.c_int, [_P, C.c_char_p, _I64]),
    "wrs_ctx_get_stat": (_I64, [_P, C.c_char_p]),
    "wrs_decoder_create": (C.c_int, [_P, C.POINTER(DecodeCfg), _P, _I32, _I32, C.POINTER(_P)]),
    "wrs_decoder_destroy": (None, [_P]),
    "wrs_decoder_reset": (C.c_int, [_P, _P]),
    "wrs_filter_apply": (C.c_int, [_P, _I32, _P, _I64, _I64, _P, _I64, _I64, _I64, _P, _P]),
    "wrs_beam_step": (C.c_int, [_P, _P, _I64, _I64, _P, _I64, _I64, _I64, _P, _P, _I64, _P, _P, _P,
                                C.POINTER(StepDebug), _P]),
    "wrs_greedy_step": (C.c_int, [_P, _P, _I64, _I64, _P, _I64, _I64, _I64, _P, _P, _I64, _P, _P,
                                  C.POINTER(StepDebug), _P]),
    "wrs_projection_create": (C.c_int, [_P, _P, _I64, _I32, _P, _P]),
    "wrs_projection_destroy": (None, [_P]),
    "wrs_project_logits": (C.c_int, [_P, _P, _I64, _I64, _P, _I64, _P]),
    "wrs_project_beam_step": (C.c_int, [_P, _P, _I64, _I64, _P, _I64, _I64, _I64, _P, _P, _I64, _P, _P, _P,
                                        C.POINTER(StepDebug), _P]),
    "wrs_project_greedy_step": (C.c_int, [_P, _P, _I64, _I64, _P, _I64, _I64, _I64, _P, _P, _I64, _P, _P,
                                          C.POINTER(StepDebug), _P]),
    "wrs_beam_finalize": (C.c_int, [_P, _P, _I64, _I64, _I64, _P, _I32, _P, _P, _P, _P, _P]),
    "wrs_rank_max_likelihood_host": (C.c_int, [_P, _P, _I32, _I32, _I32, _D, _P]),
    "wrs_kv_reorder": (C.c_int, [_P, C.POINTER(KvTensor), _I32, _I64, _P, _P]),
    "wrs_no_speech_probs": (C.c_int, [_P, _P, _I64, _I64, _P, _P]),
    "wrs_avg_logprob_host": (C.c_int, [_P, _P, _I32, _P]),
}

_lib = None


def lib() -> C.CDLL:
    """The loaded library.  Raises if it has not been built (python -m whisper_rs_b200.build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise WrsError(WRS_ERR_CUDA, f"{LIB_PATH} is missing: build it with `python -m whisper_rs_b200.build` "
                                         "(there is no CPU fallback)")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)   # AttributeError if the library does not export a declared symbol
            fn.restype, fn.argtypes = res, args
        _lib = handle
    return _lib


def check(rc: int, ctx_handle=None):
    if rc != WRS_OK:
        msg = lib().wrs_last_error(ctx_handle)
        raise WrsError(rc, msg.decode() if msg else "?")


class Context:
    """wrs_ctx: one per device."""

    def __init__(self, device: int = 0):
        self.handle = _P()
        check(lib().wrs_ctx_create(int(device), C.byref(self.handle)))
        self.device = int(device)

    def close(self):
        if getattr(self, "handle", None):
            lib().wrs_ctx_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def check(self, stream: int = 0):
        check(lib().wrs_ctx_check(self.handle, _P(stream)), self.handle)

    def set_option(self, key: str, value: int):
        check(lib().wrs_ctx_set_option(self.handle, key.encode(), int(value)), self.handle)

    def get_stat(self, key: str) -> int:
        return int(lib().wrs_ctx_get_stat(self.handle, key.encode()))


_contexts = {}


def context(device: int = 0) -> Context:
    """Process-wide default context of a device."""
    if device not in _contexts:
        _contexts[device] = Context(device)
    return _contexts[device]
