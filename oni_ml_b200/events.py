"""Python face of the event pipelines around the scorer (flow_post_lda.scala,
dns_post_lda.scala): raw rows -> features -> indices -> scored rows.  Forwards
to liboni_lda.so (csrc/events.cu); no featurisation or arithmetic happens here.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import check, ptr


def _enc(path):
    return None if path is None else path.encode()


class _Events:
    _CUTS: tuple = ()

    def __init__(self):
        self._L = _lib.load()
        self._h = C.c_void_p()

    def close(self):
        if self._h:
            self._L.oni_events_close(self._h)
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

    def info(self) -> dict:
        a, b, d, w = C.c_int64(), C.c_int64(), C.c_int32(), C.c_int32()
        check(self._L.oni_events_info(self._h, C.byref(a), C.byref(b), C.byref(d), C.byref(w)))
        return dict(lines_in=a.value, rows=b.value, num_docs=d.value, num_words=w.value)

    def cuts(self) -> dict:
        out = {}
        for which, name in enumerate(self._CUTS):
            buf = np.zeros(10)
            n = C.c_int32()
            check(self._L.oni_events_cuts(self._h, which, ptr(buf), 10, C.byref(n)))
            out[name] = buf[: n.value].copy()
        return out

    def row_text(self, row: int) -> str:
        n = self._L.oni_events_row_text(self._h, row, None, 0)
        if n < 0:
            check(1)
        buf = C.create_string_buffer(n + 1)
        self._L.oni_events_row_text(self._h, row, buf, n + 1)
        return buf.value.decode("utf-8", "replace")

    def word_counts(self, out_path: str) -> int:
        n = C.c_int64()
        check(self._L.oni_events_word_counts(self._h, out_path.encode(), C.byref(n)))
        return n.value

    def score_write(self, tol: float, out_path: str, device: int = 0) -> int:
        n = C.c_int64()
        check(self._L.oni_events_score_write(self._h, float(tol), out_path.encode(), device, C.byref(n)))
        return n.value


class FlowEvents(_Events):
    """oni_flow_events_open: flow_post_lda.scala up to the score loop."""
    _CUTS = ("time", "ibyt", "ipkt")

    def __init__(self, flow_csv: str, doc_results_csv: str | None = None, word_results_csv: str | None = None,
                 feedback_csv: str | None = None, dup_factor: int = 0):
        super().__init__()
        if feedback_csv is None:
            check(self._L.oni_flow_events_open(flow_csv.encode(), _enc(doc_results_csv), _enc(word_results_csv),
                                               C.byref(self._h)))
        else:
            check(self._L.oni_flow_events_open_feedback(flow_csv.encode(), feedback_csv.encode(), dup_factor,
                                                        _enc(doc_results_csv), _enc(word_results_csv),
                                                        C.byref(self._h)))

    def indices(self):
        n = self.info()["rows"]
        a = [np.zeros(max(n, 1), dtype=np.int32) for _ in range(4)]
        check(self._L.oni_events_indices(self._h, *(ptr(x) for x in a)))
        return tuple(x[:n] for x in a)  # src_doc, dst_doc, src_word, dst_word


class DnsEvents(_Events):
    """oni_dns_events_open: dns_post_lda.scala up to the score loop."""
    _CUTS = ("unix_tstamp", "frame_len", "subdomain_length", "subdomain_entropy", "num_periods")

    def __init__(self, dns_csv: str, top_domains_csv: str, doc_results_csv: str | None = None,
                 word_results_csv: str | None = None, feedback_csv: str | None = None, dup_factor: int = 0):
        super().__init__()
        if feedback_csv is None:
            check(self._L.oni_dns_events_open(dns_csv.encode(), top_domains_csv.encode(), _enc(doc_results_csv),
                                              _enc(word_results_csv), C.byref(self._h)))
        else:
            check(self._L.oni_dns_events_open_feedback(dns_csv.encode(), top_domains_csv.encode(),
                                                       feedback_csv.encode(), dup_factor, _enc(doc_results_csv),
                                                       _enc(word_results_csv), C.byref(self._h)))

    def indices(self):
        n = self.info()["rows"]
        d, w = np.zeros(max(n, 1), dtype=np.int32), np.zeros(max(n, 1), dtype=np.int32)
        check(self._L.oni_events_indices(self._h, ptr(d), None, ptr(w), None))
        return d[:n], w[:n]


DNS_COLUMNS = ("frame_time", "unix_tstamp", "frame_len", "ip_dst", "dns_qry_name", "dns_qry_class", "dns_qry_type",
               "dns_qry_rcode")


def dns_parquet_to_csv(file_list: str, out_csv: str, as_reference: bool = True) -> int:
    """The parquet front end of dns_pre_lda.scala:140-152 / dns_post_lda.scala:108-121: DNS_PATH is a
    comma-separated list of parquet files; rows with a null frame_len or unix_tstamp are dropped, the
    eight columns the scripts select are joined with ',' (Row.mkString: null prints as "null"), one
    line per event, ready for oni_dns_events_open.  `as_reference` keeps the scripts' loop condition
    `index > 1`, which never reads the SECOND file of the list; pass False to read every file.
    Host-side glue with no arithmetic; needs pyarrow.  Returns the number of lines written."""
    import pyarrow.parquet as pq

    files = file_list.split(",")
    use = [f for i, f in enumerate(files) if i == 0 or i > 1 or not as_reference]
    n = 0
    with open(out_csv, "w") as out:
        for path in use:
            t = pq.read_table(path, columns=list(DNS_COLUMNS))
            cols = [t.column(c).to_pylist() for c in DNS_COLUMNS]
            for row in zip(*cols):
                if row[2] is None or row[1] is None:
                    continue
                out.write(",".join("null" if v is None else str(v) for v in row) + "\n")
                n += 1
    return n
