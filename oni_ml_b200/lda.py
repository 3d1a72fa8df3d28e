"""Host-side mirror of the engine's C ABI for Python callers (tests, bench).

Names follow the reference's operations: an :class:`LdaEngine` is one rank of
``run_em`` (oni-lda-c/lda-estimate.c:112-321) living on one GPU; ``estimate``
is the whole run over host buffers; :func:`run_cli` is the ``lda`` executable.
Everything here forwards to liboni_lda.so -- no arithmetic happens in Python.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from . import _lib
from ._lib import Settings, check
from .corpus import Corpus


def read_settings(path: str) -> Settings:
    """read_settings (lda-estimate.c:331-350) through the library."""
    st = Settings()
    check(_lib.load().oni_lda_read_settings(path.encode(), C.byref(st)))
    return st


def read_corpus(path: str) -> Corpus:
    """read_data (lda-data.c:22-59) through the library's parser."""
    L = _lib.load()
    c = _lib.CorpusC()
    check(L.oni_lda_read_corpus(path.encode(), C.byref(c)))
    try:
        d, nnz = c.num_docs, c.nnz
        ptr = np.ctypeslib.as_array(c.doc_ptr, shape=(d + 1,)).copy()
        word = np.ctypeslib.as_array(c.word, shape=(max(nnz, 1),))[:nnz].copy()
        count = np.ctypeslib.as_array(c.count, shape=(max(nnz, 1),))[:nnz].copy()
        return Corpus(ptr, word, count, c.num_terms)
    finally:
        L.oni_lda_free_corpus(C.byref(c))


def lda_pre(dir_prefix: str) -> None:
    """lda_pre.py (ml_ops.sh:67) through the library: doc_wc.dat -> words.dat, doc.dat, model.dat."""
    check(_lib.load().oni_lda_pre(dir_prefix.encode()))


def _corpus_c(corpus: Corpus) -> _lib.CorpusC:
    ptr = np.ascontiguousarray(corpus.doc_ptr, dtype=np.int64)
    word = np.ascontiguousarray(corpus.word, dtype=np.int32)
    count = np.ascontiguousarray(corpus.count, dtype=np.int32)
    c = _lib.CorpusC(corpus.num_docs, corpus.num_terms, corpus.nnz,
                     ptr.ctypes.data_as(C.POINTER(C.c_int64)),
                     word.ctypes.data_as(C.POINTER(C.c_int32)),
                     count.ctypes.data_as(C.POINTER(C.c_int32)))
    c._keep = (ptr, word, count)
    return c


def write_corpus(path: str, corpus: Corpus) -> None:
    """model.dat text of a corpus (lda_pre.py:94) through the library's threaded writer."""
    check(_lib.load().oni_lda_write_corpus(path.encode(), C.byref(_corpus_c(corpus))))


def estimate(corpus: Corpus, num_topics: int, alpha: float, settings: Settings, seed: int,
             ngpus: int = 1) -> dict:
    """oni_lda_estimate: run_em with ``random`` start over host buffers."""
    L = _lib.load()
    k, v, d = num_topics, corpus.num_terms, corpus.num_docs
    cap = settings.em_max_iter + 2
    lb0 = np.zeros((k, v)); lb = np.zeros((k, v)); gam = np.zeros((d, k))
    lik = np.zeros(cap); conv = np.zeros(cap)
    dp = C.POINTER(C.c_double)
    res = _lib.ResultC(lb0.ctypes.data_as(dp), lb.ctypes.data_as(dp), gam.ctypes.data_as(dp),
                       lik.ctypes.data_as(dp), conv.ctypes.data_as(dp), cap, 0, 0.0, 0.0)
    cc = _corpus_c(corpus)
    check(L.oni_lda_estimate(C.byref(cc), k, float(alpha), C.byref(settings), seed & 0xFFFFFFFF,
                             ngpus, C.byref(res)))
    n = res.em_iters
    return dict(logbeta0=lb0, logbeta=lb, gamma=gam, alpha=res.alpha, lik=lik[:n], conv=conv[:n],
                iters=n, seconds_em=res.seconds_em)


class LdaEngine:
    """One GPU's shard of an EM run (oni_lda_ctx)."""

    def __init__(self, device: int, num_topics: int, num_terms: int, settings: Settings):
        self._L = _lib.load()
        self._h = C.c_void_p()
        self.num_topics, self.num_terms = num_topics, num_terms
        self.settings = settings
        self.num_docs = 0
        check(self._L.oni_lda_create(device, num_topics, num_terms, C.byref(settings),
                                     C.byref(self._h)))

    def close(self):
        if self._h:
            self._L.oni_lda_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- corpus ------------------------------------------------------------
    def set_corpus(self, corpus: Corpus):
        ptr = np.ascontiguousarray(corpus.doc_ptr, dtype=np.int64)
        word = np.ascontiguousarray(corpus.word, dtype=np.int32)
        count = np.ascontiguousarray(corpus.count, dtype=np.int32)
        check(self._L.oni_lda_set_corpus(self._h, corpus.num_docs, corpus.nnz, ptr.ctypes.data,
                                         word.ctypes.data, count.ctypes.data))
        self.num_docs = corpus.num_docs

    def set_corpus_device(self, doc_ptr, word, count):
        """torch tensors on this engine's device: int64 [D+1], int32 [NNZ], int32 [NNZ]."""
        d, nnz = int(doc_ptr.numel()) - 1, int(word.numel())
        check(self._L.oni_lda_set_corpus_device(self._h, d, nnz, doc_ptr.data_ptr(),
                                                word.data_ptr(), count.data_ptr()))
        self.num_docs = d

    # -- model -------------------------------------------------------------
    def init_random(self, seed: int, alpha: float):
        check(self._L.oni_lda_init_random(self._h, seed & 0xFFFFFFFF, float(alpha)))

    def set_model(self, logbeta: np.ndarray, alpha: float):
        lb = np.ascontiguousarray(logbeta, dtype=np.float64)
        assert lb.shape == (self.num_topics, self.num_terms)
        check(self._L.oni_lda_set_model(self._h, lb.ctypes.data, float(alpha)))

    # -- steps -------------------------------------------------------------
    def e_step(self):
        check(self._L.oni_lda_e_step(self._h))

    def allreduce(self):
        check(self._L.oni_lda_allreduce(self._h))

    def m_step(self):
        lik, conv = C.c_double(), C.c_double()
        check(self._L.oni_lda_m_step(self._h, C.byref(lik), C.byref(conv)))
        return lik.value, conv.value

    def em_iteration(self):
        lik, conv = C.c_double(), C.c_double()
        check(self._L.oni_lda_em_iteration(self._h, C.byref(lik), C.byref(conv)))
        return lik.value, conv.value

    def should_continue(self) -> bool:
        return bool(self._L.oni_lda_em_should_continue(self._h))

    def suffstats_ptr(self):
        p, n = C.c_void_p(), C.c_size_t()
        check(self._L.oni_lda_suffstats(self._h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def comm_init(self, unique_id: bytes, rank: int, nranks: int):
        buf = (C.c_ubyte * _lib.NCCL_ID_BYTES).from_buffer_copy(unique_id)
        check(self._L.oni_lda_comm_init(self._h, C.addressof(buf), rank, nranks))

    # -- readback ----------------------------------------------------------
    def _out(self, out, shape):
        if out is None:
            return np.empty(shape)
        if out.shape != shape or out.dtype != np.float64 or not out.flags["C_CONTIGUOUS"]:
            raise ValueError("out must be a C-contiguous float64 array of shape %r" % (shape,))
        return out

    def logbeta(self, out: np.ndarray | None = None) -> np.ndarray:
        """final.beta values [K][V]; `out` (e.g. a pinned buffer the caller reuses) is filled in place."""
        out = self._out(out, (self.num_topics, self.num_terms))
        check(self._L.oni_lda_get_logbeta(self._h, out.ctypes.data))
        return out

    def gamma(self, out: np.ndarray | None = None) -> np.ndarray:
        """final.gamma values [D][K]; `out` as for logbeta."""
        out = self._out(out, (self.num_docs, self.num_topics))
        check(self._L.oni_lda_get_gamma(self._h, out.ctypes.data))
        return out

    def alpha(self) -> float:
        a = C.c_double()
        check(self._L.oni_lda_get_alpha(self._h, C.byref(a)))
        return a.value

    def doc_likelihood(self) -> np.ndarray:
        out = np.zeros(self.num_docs)
        check(self._L.oni_lda_get_doc_likelihood(self._h, out.ctypes.data))
        return out

    def doc_iters(self) -> np.ndarray:
        out = np.zeros(self.num_docs, dtype=np.int32)
        check(self._L.oni_lda_get_doc_iters(self._h, out.ctypes.data))
        return out

    def stream_ptr(self) -> int:
        """cudaStream_t of this engine (wrap with torch.cuda.ExternalStream to time steps)."""
        s = C.c_void_p()
        check(self._L.oni_lda_stream(self._h, C.byref(s)))
        return s.value or 0

    def stats(self) -> dict:
        s = _lib.StatsC()
        check(self._L.oni_lda_get_stats(self._h, C.byref(s)))
        return {f: getattr(s, f) for f, _ in s._fields_}


def nccl_unique_id() -> bytes:
    buf = (C.c_ubyte * _lib.NCCL_ID_BYTES)()
    check(_lib.load().oni_lda_comm_unique_id(C.addressof(buf)))
    return bytes(buf)


def save_model(root: str, logbeta: np.ndarray, alpha: float):
    lb = np.ascontiguousarray(logbeta, dtype=np.float64)
    check(_lib.load().oni_lda_save_model(root.encode(), lb.shape[0], lb.shape[1], lb.ctypes.data,
                                         float(alpha)))


def save_gamma(path: str, gamma: np.ndarray):
    g = np.ascontiguousarray(gamma, dtype=np.float64)
    check(_lib.load().oni_lda_save_gamma(path.encode(), g.shape[0], g.shape[1], g.ctypes.data))


def run_cli(args: list[str], seed: int | None = None, check_rc: bool = True):
    """Run the `lda` executable (the drop-in for oni-lda-c's binary)."""
    env = dict(os.environ)
    if seed is not None:
        env["ONI_LDA_SEED"] = str(seed)
    p = subprocess.run([_lib.LDA_BIN] + list(args), env=env, capture_output=True, text=True)
    if check_rc and p.returncode != 0:
        raise RuntimeError("lda %s failed (%d): %s" % (" ".join(args), p.returncode, p.stderr))
    return p
