"""Python face of the event scorer (flow_post_lda.scala:227-245,
dns_post_lda.scala:312-326) and of the lda_post.py table shaping.  Forwards to
liboni_lda.so; no arithmetic happens here.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import check, ptr


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def score_flow(theta, pwz, src_doc, dst_doc, src_word, dst_word, tol: float, out_cap: int | None = None,
               device: int = 0, sides: bool = True, ngpus: int = 0) -> dict:
    """oni_score_flow over host arrays.  theta [D][K], pwz [V][K]; index < 0 = miss.
    ngpus > 0: oni_score_flow_multi over devices 0..ngpus-1 (`device` is then ignored)."""
    L = _lib.load()
    theta, pwz = _f64(theta), _f64(pwz)
    sd, dd, sw, dw = map(_i32, (src_doc, dst_doc, src_word, dst_word))
    n = sd.shape[0]
    cap = n if out_cap is None else int(out_cap)
    idx = np.zeros(max(cap, 1), dtype=np.int64); key = np.zeros(max(cap, 1))
    sa = np.zeros(max(cap, 1)) if sides else None
    sb = np.zeros(max(cap, 1)) if sides else None
    total = C.c_int64()
    fn, first = (L.oni_score_flow_multi, ngpus) if ngpus > 0 else (L.oni_score_flow, device)
    check(fn(first, theta.shape[1], theta.shape[0], pwz.shape[0], ptr(theta), ptr(pwz),
             n, ptr(sd), ptr(dd), ptr(sw), ptr(dw), float(tol), cap, ptr(idx), ptr(key),
             ptr(sa), ptr(sb), C.byref(total)))
    m = min(cap, total.value)
    out = dict(n=total.value, idx=idx[:m], key=key[:m])
    if sides:
        out.update(src=sa[:m], dst=sb[:m])
    return out


def score_dns(theta, pwz, doc, word, tol: float, out_cap: int | None = None, device: int = 0,
              ngpus: int = 0) -> dict:
    L = _lib.load()
    theta, pwz = _f64(theta), _f64(pwz)
    d, w = _i32(doc), _i32(word)
    n = d.shape[0]
    cap = n if out_cap is None else int(out_cap)
    idx = np.zeros(max(cap, 1), dtype=np.int64); key = np.zeros(max(cap, 1))
    total = C.c_int64()
    fn, first = (L.oni_score_dns_multi, ngpus) if ngpus > 0 else (L.oni_score_dns, device)
    check(fn(first, theta.shape[1], theta.shape[0], pwz.shape[0], ptr(theta), ptr(pwz), n,
             ptr(d), ptr(w), float(tol), cap, ptr(idx), ptr(key), C.byref(total)))
    m = min(cap, total.value)
    return dict(n=total.value, idx=idx[:m], key=key[:m])


class ScoreContext:
    """oni_score_ctx: tables resident on one GPU; events arrive as device tensors."""

    def __init__(self, theta, pwz, device: int = 0):
        self._L = _lib.load()
        theta, pwz = _f64(theta), _f64(pwz)
        self._h = C.c_void_p()
        check(self._L.oni_score_create(device, theta.shape[1], theta.shape[0], pwz.shape[0],
                                       ptr(theta), ptr(pwz), C.byref(self._h)))

    def close(self):
        if self._h:
            self._L.oni_score_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def begin(self):
        check(self._L.oni_score_begin(self._h))

    def score_flow(self, src_doc, dst_doc, src_word, dst_word, tol: float, idx_base: int = 0):
        check(self._L.oni_score_events_device(self._h, 1, int(src_doc.numel()), idx_base, ptr(src_doc),
                                              ptr(dst_doc), ptr(src_word), ptr(dst_word), float(tol)))

    def score_dns(self, doc, word, tol: float, idx_base: int = 0):
        check(self._L.oni_score_events_device(self._h, 0, int(doc.numel()), idx_base, ptr(doc), None,
                                              ptr(word), None, float(tol)))

    def score_flow_host(self, src_doc, dst_doc, src_word, dst_word, tol: float, idx_base: int = 0):
        """Events in host arrays (numpy int32; pinned memory travels at PCIe speed): uploaded in
        chunks while the previous chunk is scored (oni_score_events_host)."""
        a = [np.ascontiguousarray(x, dtype=np.int32) for x in (src_doc, dst_doc, src_word, dst_word)]
        check(self._L.oni_score_events_host(self._h, 1, int(a[0].shape[0]), idx_base, ptr(a[0]), ptr(a[1]),
                                            ptr(a[2]), ptr(a[3]), float(tol)))

    def score_dns_host(self, doc, word, tol: float, idx_base: int = 0):
        a = [np.ascontiguousarray(x, dtype=np.int32) for x in (doc, word)]
        check(self._L.oni_score_events_host(self._h, 0, int(a[0].shape[0]), idx_base, ptr(a[0]), None,
                                            ptr(a[1]), None, float(tol)))

    def rank(self) -> int:
        n = C.c_int64()
        check(self._L.oni_score_rank(self._h, C.byref(n)))
        return n.value

    def fetch(self, cap: int):
        idx = np.zeros(max(cap, 1), dtype=np.int64); key = np.zeros(max(cap, 1))
        check(self._L.oni_score_fetch(self._h, cap, ptr(idx), ptr(key)))
        return idx[:cap], key[:cap]

    def stats(self) -> dict:
        a, b, n = C.c_double(), C.c_double(), C.c_int64()
        check(self._L.oni_score_get_stats(self._h, C.byref(a), C.byref(b), C.byref(n)))
        return dict(ms_score=a.value, ms_sort=b.value, kernel_launches=n.value)


def lda_post(gamma, logbeta, text_digits: int = 12, device: int = 0):
    """oni_lda_post: theta [D][K] and pwz [V][K] from final.gamma / final.beta values."""
    L = _lib.load()
    gamma, logbeta = _f64(gamma), _f64(logbeta)
    d, k = gamma.shape
    v = logbeta.shape[1]
    theta = np.zeros((d, k)); pwz = np.zeros((v, k))
    check(L.oni_lda_post(device, d, k, v, ptr(gamma), ptr(logbeta), text_digits, ptr(theta), ptr(pwz)))
    return theta, pwz


def lda_post_files(dir_prefix: str, device: int = 0) -> None:
    """oni_lda_post_files: lda_post.py as a whole (doc_results.csv, word_results.csv)."""
    check(_lib.load().oni_lda_post_files(dir_prefix.encode(), device))
