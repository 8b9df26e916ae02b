"""ctypes binding of libb200vec.so — the C-ABI declared in include/b200vec.h.

The product path has NO CPU fallback: if the shared library is missing, or no B200-class
CUDA device is visible, importing the engine raises with instructions instead of silently
computing on the host.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200VEC_LIB") or os.path.join(_HERE, "csrc", "libb200vec.so")   # B200VEC_LIB: A/B builds

MAX_PACKETS = 32
MAX_SLOTS = 64
MAX_RANKS = 8
BUILD_STORE, BUILD_MAXONLY, BUILD_SCALE = 0, 1, 2
ABI_VERSION = 2


class PktDesc(C.Structure):
    """b200vec_pkt_desc (include/b200vec.h)."""
    _fields_ = [("start", C.c_int64), ("period", C.c_int64), ("count", C.c_int64),
                ("slot", C.c_int32), ("reserved", C.c_int32)]


class BatchPkt(C.Structure):
    """b200vec_batch_pkt (include/b200vec.h)."""
    _fields_ = [("start", C.c_int64), ("period", C.c_int64), ("count", C.c_int64),
                ("src_slot", C.c_int32), ("flags", C.c_int32), ("freq_hz", C.c_double)]


class B200VecError(RuntimeError):
    pass


_SIGNATURES = {
    "b200vec_abi_version": (C.c_int, []),
    "b200vec_last_error": (C.c_char_p, []),
    "b200vec_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "b200vec_destroy": (C.c_int, [C.c_void_p]),
    "b200vec_bank_upload": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_int, C.c_void_p]),
    "b200vec_bank_resample": (C.c_int, [C.c_void_p, C.c_int, C.c_int64, C.c_void_p]),
    "b200vec_bank_shift": (C.c_int, [C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_void_p]),
    "b200vec_bank_attach": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_void_p]),
    "b200vec_bank_len": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_int64)]),
    "b200vec_bank_download": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_void_p]),
    "b200vec_build": (C.c_int, [C.c_void_p, C.POINTER(PktDesc), C.c_int, C.c_int64, C.c_int64,
                                C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "b200vec_build_norm": (C.c_int, [C.c_void_p, C.POINTER(PktDesc), C.c_int, C.c_int64, C.c_int64,
                                     C.c_void_p, C.POINTER(C.c_int64), C.c_int, C.c_void_p, C.c_void_p]),
    "b200vec_batch_prepare": (C.c_int, [C.c_void_p, C.POINTER(BatchPkt), C.POINTER(C.c_int32), C.c_int, C.c_double,
                                        C.c_void_p]),
    "b200vec_batch_build": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p]),
    "b200vec_peer_status": (C.c_int, [C.c_void_p, C.POINTER(C.c_int), C.c_void_p]),
    "b200vec_select_begin": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.POINTER(C.c_int64), C.c_void_p]),
    "b200vec_select_finish": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(C.c_int64), C.c_int,
                                        C.POINTER(C.c_float), C.c_void_p]),
    "b200vec_decimate": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]),
    "b200vec_stft": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int,
                               C.c_int, C.c_int, C.c_void_p, C.c_float, C.c_int64, C.c_int64,
                               C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "b200vec_stft_f64": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int,
                                   C.c_int, C.c_int, C.c_void_p, C.c_double, C.c_int64, C.c_int64,
                                   C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "b200vec_build_stft1024": (C.c_int, [C.c_void_p, C.POINTER(PktDesc), C.c_int, C.c_int64, C.c_int64,
                                         C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_float,
                                         C.c_int64, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "b200vec_build_stft1024_fm": (C.c_int, [C.c_void_p, C.POINTER(PktDesc), C.c_int, C.c_int64, C.c_int64,
                                            C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_float,
                                            C.c_int64, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "b200vec_build_stft1024_fm_norm": (C.c_int, [C.c_void_p, C.POINTER(PktDesc), C.c_int, C.c_int64, C.c_int64,
                                                 C.c_void_p, C.POINTER(C.c_int64), C.c_int, C.c_void_p,
                                                 C.c_void_p, C.c_float, C.c_int64, C.c_int64, C.c_void_p,
                                                 C.c_int64, C.c_void_p, C.c_void_p]),
    "b200vec_stft1024_fm": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_float,
                                      C.c_int64, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "b200vec_build_stft1024_fm_norm_peer": (C.c_int, [C.c_void_p, C.POINTER(PktDesc), C.c_int, C.c_int64, C.c_int64,
                                                      C.c_void_p, C.POINTER(C.c_int64), C.c_int,
                                                      C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_uint32, C.c_void_p,
                                                      C.c_void_p, C.c_float, C.c_int64, C.c_int64, C.c_void_p,
                                                      C.c_int64, C.c_void_p, C.c_void_p]),
    "b200vec_db": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_void_p, C.c_void_p]),
    "b200vec_replicate_frames": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int64,
                                           C.c_void_p]),
    "b200vec_normalize_db": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_double, C.c_double, C.c_void_p, C.c_void_p,
                                       C.c_void_p, C.c_void_p]),
    "b200vec_select_f64": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "b200vec_db_values": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_void_p, C.c_void_p, C.c_int,
                                    C.c_void_p, C.c_void_p]),
    "b200vec_select_count": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int,
                                       C.POINTER(C.c_int64), C.c_void_p]),
    "b200vec_select_ranks": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int,
                                       C.POINTER(C.c_int64), C.c_int, C.POINTER(C.c_float),
                                       C.c_void_p]),
    "b200vec_wv_quantize": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_void_p, C.c_void_p]),
    "b200vec_power_stats": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.POINTER(C.c_double), C.c_void_p]),
    "b200vec_power_stats_static": (C.c_int, [C.c_int, C.c_void_p, C.c_int64, C.c_float, C.c_void_p, C.c_void_p]),
    "b200vec_box_energy": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200vec_threshold_bounds": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_double, C.POINTER(C.c_int64), C.c_void_p]),
    "b200vec_median2_rows": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64,
                                       C.c_void_p, C.c_int64, C.c_int64, C.c_void_p]),
    "b200vec_split_planes": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
}

EXPORTED_SYMBOLS = tuple(_SIGNATURES)
_lib = None


def load():
    """dlopen libb200vec.so (once) and attach signatures.  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise B200VecError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; "
            "g.build()'` (or `make -C vector_b200/csrc`).  vector_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the .so does not export it
        fn.restype = res
        fn.argtypes = args
    if lib.b200vec_abi_version() != ABI_VERSION:
        raise B200VecError("libb200vec.so ABI version mismatch; rebuild it")
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        msg = load().b200vec_last_error()
        raise B200VecError(msg.decode() if msg else f"libb200vec error {rc}")
