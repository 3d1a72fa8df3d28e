"""ctypes binding of liboni_lda.so (include/oni_lda.h).

The library is the product; this module only declares its C ABI.  There is no
fallback: if the shared object is missing the import fails with instructions to
build it, and compute calls fail inside the library when no CUDA device exists.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# ONI_LDA_LIB selects another build of the same library (kernel tuning experiments)
LIB_PATH = os.environ.get("ONI_LDA_LIB") or os.path.join(_HERE, "liboni_lda.so")
LDA_BIN = os.path.join(_HERE, "bin", "lda")

NCCL_ID_BYTES = 128


class OniError(RuntimeError):
    """A non-zero oni_status from the library."""

    def __init__(self, code: int, msg: str):
        super().__init__("oni status %d: %s" % (code, msg))
        self.code = code


class Settings(C.Structure):
    """oni_lda_settings: the five values of settings.txt
    (oni-lda-c/lda-estimate.c:331-350); thresholds are floats as in the reference."""
    _fields_ = [("var_max_iter", C.c_int32), ("var_converged", C.c_float),
                ("em_max_iter", C.c_int32), ("em_converged", C.c_float),
                ("estimate_alpha", C.c_int32)]

    @classmethod
    def shipped(cls) -> "Settings":
        """oni-lda-c/settings.txt as shipped."""
        return cls(20, 1e-6, 20, 1e-6, 1)


class CorpusC(C.Structure):
    _fields_ = [("num_docs", C.c_int32), ("num_terms", C.c_int32), ("nnz", C.c_int64),
                ("doc_ptr", C.POINTER(C.c_int64)), ("word", C.POINTER(C.c_int32)),
                ("count", C.POINTER(C.c_int32))]


class ResultC(C.Structure):
    _fields_ = [("logbeta_init", C.POINTER(C.c_double)), ("logbeta", C.POINTER(C.c_double)),
                ("gamma", C.POINTER(C.c_double)), ("likelihood", C.POINTER(C.c_double)),
                ("converged", C.POINTER(C.c_double)), ("likelihood_cap", C.c_int32),
                ("em_iters", C.c_int32), ("alpha", C.c_double), ("seconds_em", C.c_double)]


class StatsC(C.Structure):
    _fields_ = [("ms_estep_kernel", C.c_double), ("ms_suffstats", C.c_double),
                ("ms_mstep", C.c_double), ("word_visits", C.c_int64),
                ("kernel_launches", C.c_int64),
                ("ms_end_exclusive", C.c_double), ("ms_end_long", C.c_double), ("ms_end_main", C.c_double),
                ("docs_exclusive", C.c_int32), ("docs_long", C.c_int32), ("longest_doc_pairs", C.c_int64)]


_vp = C.c_void_p
_i32, _i64, _f64 = C.c_int32, C.c_int64, C.c_double

# name -> (restype, argtypes); every symbol include/oni_lda.h declares
SIGNATURES = {
    "oni_last_error": (C.c_char_p, []),
    "oni_version": (C.c_char_p, []),
    "oni_lda_read_settings": (C.c_int, [C.c_char_p, C.POINTER(Settings)]),
    "oni_lda_read_corpus": (C.c_int, [C.c_char_p, C.POINTER(CorpusC)]),
    "oni_lda_free_corpus": (None, [C.POINTER(CorpusC)]),
    "oni_lda_write_corpus": (C.c_int, [C.c_char_p, C.POINTER(CorpusC)]),
    "oni_lda_pre": (C.c_int, [C.c_char_p]),
    "oni_lda_main": (C.c_int, [C.c_int, C.POINTER(C.c_char_p)]),
    "oni_lda_estimate": (C.c_int, [C.POINTER(CorpusC), _i32, _f64, C.POINTER(Settings), C.c_uint32,
                                   _i32, C.POINTER(ResultC)]),
    "oni_lda_create": (C.c_int, [_i32, _i32, _i32, C.POINTER(Settings), C.POINTER(_vp)]),
    "oni_lda_destroy": (None, [_vp]),
    "oni_lda_set_corpus": (C.c_int, [_vp, _i32, _i64, _vp, _vp, _vp]),
    "oni_lda_set_corpus_device": (C.c_int, [_vp, _i32, _i64, _vp, _vp, _vp]),
    "oni_lda_init_random": (C.c_int, [_vp, C.c_uint32, _f64]),
    "oni_lda_set_model": (C.c_int, [_vp, _vp, _f64]),
    "oni_lda_e_step": (C.c_int, [_vp]),
    "oni_lda_suffstats": (C.c_int, [_vp, C.POINTER(_vp), C.POINTER(C.c_size_t)]),
    "oni_lda_comm_unique_id": (C.c_int, [_vp]),
    "oni_lda_comm_init": (C.c_int, [_vp, _vp, _i32, _i32]),
    "oni_lda_allreduce": (C.c_int, [_vp]),
    "oni_lda_m_step": (C.c_int, [_vp, C.POINTER(_f64), C.POINTER(_f64)]),
    "oni_lda_em_iteration": (C.c_int, [_vp, C.POINTER(_f64), C.POINTER(_f64)]),
    "oni_lda_em_should_continue": (C.c_int, [_vp]),
    "oni_lda_get_logbeta": (C.c_int, [_vp, _vp]),
    "oni_lda_get_gamma": (C.c_int, [_vp, _vp]),
    "oni_lda_get_alpha": (C.c_int, [_vp, C.POINTER(_f64)]),
    "oni_lda_get_doc_likelihood": (C.c_int, [_vp, _vp]),
    "oni_lda_get_doc_iters": (C.c_int, [_vp, _vp]),
    "oni_lda_get_stats": (C.c_int, [_vp, C.POINTER(StatsC)]),
    "oni_lda_stream": (C.c_int, [_vp, C.POINTER(_vp)]),
    "oni_lda_save_model": (C.c_int, [C.c_char_p, _i32, _i32, _vp, _f64]),
    "oni_lda_save_gamma": (C.c_int, [C.c_char_p, _i32, _i32, _vp]),
    "oni_score_flow": (C.c_int, [_i32, _i32, _i32, _i32, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _f64,
                                 _i64, _vp, _vp, _vp, _vp, C.POINTER(_i64)]),
    "oni_score_dns": (C.c_int, [_i32, _i32, _i32, _i32, _vp, _vp, _i64, _vp, _vp, _f64, _i64, _vp,
                                _vp, C.POINTER(_i64)]),
    "oni_score_flow_multi": (C.c_int, [_i32, _i32, _i32, _i32, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _f64,
                                 _i64, _vp, _vp, _vp, _vp, C.POINTER(_i64)]),
    "oni_score_dns_multi": (C.c_int, [_i32, _i32, _i32, _i32, _vp, _vp, _i64, _vp, _vp, _f64, _i64, _vp,
                                _vp, C.POINTER(_i64)]),
    "oni_score_create": (C.c_int, [_i32, _i32, _i32, _i32, _vp, _vp, C.POINTER(_vp)]),
    "oni_score_destroy": (None, [_vp]),
    "oni_score_begin": (C.c_int, [_vp]),
    "oni_score_events_device": (C.c_int, [_vp, _i32, _i64, _i64, _vp, _vp, _vp, _vp, _f64]),
    "oni_score_rank": (C.c_int, [_vp, C.POINTER(_i64)]),
    "oni_score_fetch": (C.c_int, [_vp, _i64, _vp, _vp]),
    "oni_score_fetch_sides": (C.c_int, [_vp, _i32, _i64, _i64, _vp, _vp, _vp, _vp, _vp, _vp]),
    "oni_score_events_host": (C.c_int, [_vp, _i32, _i64, _i64, _vp, _vp, _vp, _vp, _f64]),
    "oni_score_fetch_sides_host": (C.c_int, [_vp, _i32, _i64, _i64, _vp, _vp, _vp, _vp, _vp, _vp]),
    "oni_score_get_stats": (C.c_int, [_vp, C.POINTER(_f64), C.POINTER(_f64), C.POINTER(_i64)]),
    "oni_score_flow_noalloc": (C.c_int, [_i32, _i32, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _f64, _i64,
                                         _vp, _vp, C.POINTER(_i64), _vp]),
    "oni_lda_post_files": (C.c_int, [C.c_char_p, _i32]),
    "oni_flow_events_open": (C.c_int, [C.c_char_p, C.c_char_p, C.c_char_p, C.POINTER(_vp)]),
    "oni_dns_events_open": (C.c_int, [C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.POINTER(_vp)]),
    "oni_flow_events_open_feedback": (C.c_int, [C.c_char_p, C.c_char_p, _i32, C.c_char_p, C.c_char_p, C.POINTER(_vp)]),
    "oni_dns_events_open_feedback": (C.c_int, [C.c_char_p, C.c_char_p, C.c_char_p, _i32, C.c_char_p, C.c_char_p,
                                               C.POINTER(_vp)]),
    "oni_events_close": (None, [_vp]),
    "oni_events_info": (C.c_int, [_vp, C.POINTER(_i64), C.POINTER(_i64), C.POINTER(_i32), C.POINTER(_i32)]),
    "oni_events_cuts": (C.c_int, [_vp, _i32, _vp, _i32, C.POINTER(_i32)]),
    "oni_events_indices": (C.c_int, [_vp, _vp, _vp, _vp, _vp]),
    "oni_events_row_text": (_i64, [_vp, _i64, C.c_char_p, _i64]),
    "oni_events_word_counts": (C.c_int, [_vp, C.c_char_p, C.POINTER(_i64)]),
    "oni_events_score_write": (C.c_int, [_vp, _f64, C.c_char_p, _i32, C.POINTER(_i64)]),
    "oni_lda_post": (C.c_int, [_i32, _i32, _i32, _i32, _vp, _vp, _i32, _vp, _vp]),
}

_lib = None


def load() -> C.CDLL:
    """dlopen the engine and declare every entry point.  Fails loudly."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "%s is missing: build it with `make -C oni_ml_b200/csrc -j` (or "
            "`python -c 'import __graft_entry__ as g; g.build()'`); there is no CPU or "
            "pure-Python path" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != 0:
        raise OniError(rc, load().oni_last_error().decode("utf-8", "replace"))


def ptr(a) -> int | None:
    """Address of a numpy array (host) or of a torch tensor (host or device)."""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        assert a.flags["C_CONTIGUOUS"]
        return a.ctypes.data
    return a.data_ptr()  # torch tensor
