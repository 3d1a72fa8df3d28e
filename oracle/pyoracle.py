"""ctypes bindings for the oracle (oracle/_build/liboni_oracle.so).

TEST INFRASTRUCTURE ONLY: importable from tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs, never from oni_ml_b200.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboni_oracle.so")
REF_BIN = os.path.join(_HERE, "_ref", "lda_ref")

_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")


class Settings(C.Structure):
    """The five values of settings.txt (oni-lda-c/lda-estimate.c:331-350)."""
    _fields_ = [("var_max_iter", C.c_int), ("var_converged", C.c_float),
                ("em_max_iter", C.c_int), ("em_converged", C.c_float),
                ("estimate_alpha", C.c_int)]


def build() -> None:
    if not os.path.exists(_SO) or (os.path.getmtime(_SO) <
                                   max(os.path.getmtime(os.path.join(_HERE, f))
                                       for f in ("lda_oracle.c", "score_oracle.c"))):
        subprocess.run(["make", "-C", _HERE, "_build/liboni_oracle.so"], check=True,
                       stdout=subprocess.DEVNULL)


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        L.orc_mt_seed.argtypes = [C.c_uint32]
        L.orc_mt_next.restype = C.c_uint32
        L.orc_myrand.restype = C.c_double
        for name in ("orc_digamma", "orc_trigamma"):
            getattr(L, name).argtypes = [C.c_double]
            getattr(L, name).restype = C.c_double
        L.orc_log_sum.argtypes = [C.c_double, C.c_double]
        L.orc_log_sum.restype = C.c_double
        L.orc_opt_alpha.argtypes = [C.c_double, C.c_int, C.c_int]
        L.orc_opt_alpha.restype = C.c_double
        L.orc_mle_beta.argtypes = [C.c_int, C.c_int, _f64p, _f64p, _f64p]
        L.orc_random_init_ss.argtypes = [C.c_int, C.c_int, _f64p, _f64p]
        L.orc_doc_inference.argtypes = [C.c_int, _i32p, _i32p, C.c_int, C.c_int, C.c_int,
                                        C.c_double, _f64p, C.c_int, C.c_double, _f64p, _f64p,
                                        C.POINTER(C.c_int)]
        L.orc_doc_inference.restype = C.c_double
        L.orc_e_step.argtypes = [C.c_int, _i64p, _i32p, _i32p, C.c_int, C.c_int, C.c_int,
                                 C.c_int, C.c_double, _f64p, C.c_int, C.c_double, _f64p, _f64p,
                                 _f64p, C.POINTER(C.c_double), C.POINTER(C.c_int),
                                 C.POINTER(C.c_int64)]
        L.orc_e_step.restype = C.c_double
        L.orc_run_em.argtypes = [C.c_int, _i64p, _i32p, _i32p, C.c_int, C.c_int, C.c_double,
                                 C.POINTER(Settings), C.c_uint32, C.c_int, _f64p, _f64p, _f64p,
                                 C.POINTER(C.c_double), _f64p, _f64p, C.c_int]
        L.orc_run_em.restype = C.c_int
        L.orc_text_roundtrip.argtypes = [C.c_double, C.c_char_p]
        L.orc_text_roundtrip.restype = C.c_double
        L.orc_theta_from_gamma.argtypes = [C.c_int, C.c_int, _f64p, _f64p]
        L.orc_pwz_from_beta.argtypes = [C.c_int, C.c_int, _f64p, _f64p]
        L.orc_theta_from_gamma_raw.argtypes = [C.c_int, C.c_int, _f64p, _f64p]
        L.orc_pwz_from_beta_raw.argtypes = [C.c_int, C.c_int, _f64p, _f64p]
        L.orc_score_flow.argtypes = [C.c_int64, C.c_int, _f64p, _f64p, _i32p, _i32p, _i32p,
                                     _i32p, _f64p, _f64p]
        L.orc_score_dns.argtypes = [C.c_int64, C.c_int, _f64p, _f64p, _i32p, _i32p, _f64p]
        L.orc_rank_below.argtypes = [C.c_int64, _f64p, C.c_void_p, C.c_double, C.c_int64,
                                     _i64p, _f64p]
        L.orc_rank_below.restype = C.c_int64
        _lib = L
    return _lib


def parse_settings(text: str) -> Settings:
    """settings.txt -> Settings, the way read_settings does (positional, float
    thresholds, anything but the literal 'fixed' estimates alpha)."""
    lines = [ln.split() for ln in text.strip().splitlines()]
    return Settings(int(lines[0][-1]), float(lines[1][-1]), int(lines[2][-1]),
                    float(lines[3][-1]), 0 if lines[4][-1] == "fixed" else 1)


def run_em(corpus, k: int, alpha: float, st: Settings, seed: int, nproc: int = 1):
    """Full EM of the restatement.  Returns a dict with logbeta0/logbeta [K][V],
    gamma [D][K], alpha, lik, conv (per EM iteration)."""
    L = lib()
    D, V = corpus.num_docs, corpus.num_terms
    cap = st.em_max_iter + 2
    lb0 = np.zeros((k, V)); lb = np.zeros((k, V)); gam = np.zeros((D, k))
    lik = np.zeros(cap); conv = np.zeros(cap)
    a = C.c_double(0)
    n = L.orc_run_em(D, corpus.doc_ptr, corpus.word, corpus.count, k, V,
                     float(np.float32(alpha)),  # atof into a float, lda-estimate.c:425
                     C.byref(st), seed & 0xFFFFFFFF, nproc, lb0, lb, gam, C.byref(a), lik,
                     conv, cap)
    return dict(logbeta0=lb0, logbeta=lb, gamma=gam, alpha=a.value, lik=lik[:n], conv=conv[:n],
                iters=n)


def e_step(corpus, k: int, alpha: float, logbeta: np.ndarray, var_max_iter: int,
           var_converged: float):
    """One E-step over the whole corpus.  logbeta [K][V].  Returns dict with
    gamma [D][K], class_word [K][V], class_total [K], alpha_ss, lik, word_visits."""
    L = lib()
    D, V = corpus.num_docs, corpus.num_terms
    gam = np.zeros((D, k)); cw = np.zeros((k, V)); ct = np.zeros(k)
    ass = C.c_double(0); nd = C.c_int(0); wv = C.c_int64(0)
    lik = L.orc_e_step(D, corpus.doc_ptr, corpus.word, corpus.count, 0, 1, k, V, alpha,
                       np.ascontiguousarray(logbeta), var_max_iter,
                       float(np.float32(var_converged)), gam, cw, ct, C.byref(ass),
                       C.byref(nd), C.byref(wv))
    return dict(gamma=gam, class_word=cw, class_total=ct, alpha_ss=ass.value, lik=lik,
                num_docs=nd.value, word_visits=wv.value)
