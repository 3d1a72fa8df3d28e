"""Featurisation of raw events (csrc/events.cu) against the Python restatement of the Spark
scripts (oracle/events_oracle.py).  Host work only: runs without a GPU."""
import ctypes as C
import os
import random
import struct
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import events_data  # noqa: E402
import events_oracle as eo  # noqa: E402

from oni_ml_b200 import _lib  # noqa: E402
from oni_ml_b200.events import DnsEvents, FlowEvents  # noqa: E402


def test_double_to_string_matches_the_restatement():
    L = _lib.load()
    fn = L.oni_internal_java_double_str
    fn.restype = C.c_int
    fn.argtypes = [C.c_double, C.c_char_p]
    buf = C.create_string_buffer(64)
    rnd = random.Random(3)
    vals = [0.0, -0.0, 1.0, 80.0, 111111.0, 333333.0, 65535.0, 0.05, 0.1, 1e-3, 9.999e-4, 1e7, 9999999.0, 1.5e-20,
            0.018333333333333333, 23.999722222222225, 1e21, 1e22, 1e23, 5e-324, 1.7976931348623157e308, 2.0 ** -44,
            float("inf"), float("-inf"), float("nan"), 123456789012345680.0, 4.35, 0.3, 2.675, 100.0, 1234.5]
    vals += [rnd.random() for _ in range(2000)]
    vals += [rnd.random() * 10 ** rnd.randint(-30, 30) for _ in range(2000)]
    vals += [struct.unpack("<d", struct.pack("<Q", rnd.getrandbits(63)))[0] for _ in range(2000)]
    for v in vals:
        n = fn(v, buf)
        got = buf.value.decode()
        assert n == len(got)
        assert got == eo.jdouble(v), repr(v)
    # spot values of Double.toString known from the JDK documentation's rules
    assert eo.jdouble(80.0) == "80.0" and eo.jdouble(1e7) == "1.0E7" and eo.jdouble(1e-3) == "0.001"
    assert eo.jdouble(9.999e-4) == "9.999E-4" and eo.jdouble(0.05) == "0.05"


def _flow_case(tmp_path, n, seed):
    flow, docs, words = events_data.make_flow(str(tmp_path), n=n, seed=seed)
    text = open(flow).read()
    first = eo.flow_post(text, open(docs).read(), "", 1.0)
    vocab = []
    for r in first["rows"]:
        f = r.split(",")
        vocab += [f[33], f[34]]
    events_data.write_word_results(words, vocab)
    return flow, docs, words


def test_flow_featurisation(tmp_path):
    flow, docs, words = _flow_case(tmp_path, 4000, 7)
    ref = eo.flow_post(open(flow).read(), open(docs).read(), open(words).read(), 1.0)
    with FlowEvents(flow, docs, words) as ev:
        info = ev.info()
        assert info["rows"] == len(ref["rows"]) and info["rows"] > 3500
        assert info["lines_in"] == len(eo.text_lines(open(flow).read()))
        cuts = ev.cuts()
        for name in ("time", "ibyt", "ipkt"):
            assert cuts[name].tolist() == ref["cuts"][name], name
        for i in range(info["rows"]):
            assert ev.row_text(i) == ref["rows"][i], i
        # indices: first-appearance row of the key in the result files, -1 on a miss
        tdocs = {}
        for ln in eo.text_lines(open(docs).read()):
            tdocs.setdefault(ln.split(",")[0], len(tdocs))
        twords = {}
        for ln in eo.text_lines(open(words).read()):
            twords.setdefault(ln.split(",")[0], len(twords))
        sd, dd, sw, dw = ev.indices()
        hits = 0
        for i, r in enumerate(ref["rows"]):
            f = r.split(",")
            assert sd[i] == tdocs.get(f[8], -1) and dd[i] == tdocs.get(f[9], -1)
            assert sw[i] == twords.get(f[33], -1) and dw[i] == twords.get(f[34], -1)
            hits += (sw[i] >= 0) + (dw[i] >= 0)
        assert 0 < hits < 2 * len(ref["rows"])  # both hits and misses (20-byte names, dropped words) occur
        assert info["num_docs"] == len(tdocs) and info["num_words"] == len(twords)


def test_flow_tiny_and_degenerate_inputs(tmp_path):
    docs = tmp_path / "d.csv"
    words = tmp_path / "w.csv"
    docs.write_text("1.1.1.1," + " ".join(["0.05"] * 20) + "\n")
    words.write_text("80.0_0.0_0.0_0.0," + " ".join(["0.05"] * 20) + "\n")
    # header only / empty file: no rows, all cuts zero
    for body in ("", events_data.FLOW_HEADER + "\n"):
        p = tmp_path / "f.csv"
        p.write_text(body)
        with FlowEvents(str(p), str(docs), str(words)) as ev:
            assert ev.info()["rows"] == 0
            assert all((v == 0).all() for v in ev.cuts().values())
            assert ev.score_write(1.0, str(tmp_path / "none.csv")) == 0 and (tmp_path / "none.csv").read_text() == ""
    # one row: its own values sit at ecdf 1.0, above every quantile -> cuts stay 0, bins count cuts below the value
    rnd = random.Random(1)
    row = events_data.flow_row(rnd, ["1.1.1.1", "2.2.2.2"])
    p = tmp_path / "one.csv"
    p.write_text(events_data.FLOW_HEADER + "\n" + ",".join(row) + "\n")
    ref = eo.flow_post(p.read_text(), docs.read_text(), words.read_text(), 1.0)
    with FlowEvents(str(p), str(docs), str(words)) as ev:
        assert ev.info()["rows"] == 1 and ev.row_text(0) == ref["rows"][0]
    # a field that is not a number: the script dies, the library reports it
    bad = list(row)
    bad[16] = "many"
    p.write_text(events_data.FLOW_HEADER + "\n" + ",".join(bad) + "\n")
    with pytest.raises(_lib.OniError):
        FlowEvents(str(p), str(docs), str(words))
    with pytest.raises(_lib.OniError):
        FlowEvents(str(tmp_path / "missing.csv"), str(docs), str(words))
    # a result table with fewer than 20 numbers per line cannot feed the 0..19 loop
    words.write_text("w,0.5 0.5\n")
    p.write_text(events_data.FLOW_HEADER + "\n" + ",".join(row) + "\n")
    with pytest.raises(_lib.OniError):
        FlowEvents(str(p), str(docs), str(words))


def _dns_case(tmp_path, n, seed):
    dns, top, docs, words = events_data.make_dns(str(tmp_path), n=n, seed=seed)
    first = eo.dns_post(open(dns).read(), open(top).read(), open(docs).read(), "", 1.0)
    events_data.write_word_results(words, [r.split(",")[14] for r in first["rows"]], cut20=False)
    return dns, top, docs, words


def test_dns_featurisation(tmp_path):
    dns, top, docs, words = _dns_case(tmp_path, 4000, 13)
    ref = eo.dns_post(open(dns).read(), open(top).read(), open(docs).read(), open(words).read(), 1.0)
    with DnsEvents(dns, top, docs, words) as ev:
        info = ev.info()
        assert info["rows"] == len(ref["rows"]) and info["rows"] > 3800
        cuts = ev.cuts()
        for name, want in ref["cuts"].items():
            assert cuts[name].tolist() == want, name
        for i in range(info["rows"]):
            assert ev.row_text(i) == ref["rows"][i], i
        tdocs, twords = {}, {}
        for ln in eo.text_lines(open(docs).read()):
            tdocs.setdefault(ln.split(",")[0], len(tdocs))
        for ln in eo.text_lines(open(words).read()):
            twords.setdefault(ln.split(",")[0], len(twords))
        d, w = ev.indices()
        for i, r in enumerate(ref["rows"]):
            f = r.split(",")
            assert d[i] == tdocs.get(f[3], -1) and w[i] == twords.get(f[14], -1)
        assert (w >= 0).any() and (w < 0).any() and (d >= 0).any() and (d < 0).any()


def test_cuts_with_many_threads_match_one_thread(tmp_path, monkeypatch):
    """The sorted-slices-then-merge path (above 2^20 samples) gives the same cut vectors."""
    rnd = random.Random(5)
    ips = ["1.1.1.%d" % i for i in range(8)]
    docs = tmp_path / "d.csv"
    words = tmp_path / "w.csv"
    docs.write_text("1.1.1.1," + " ".join(["0.05"] * 20) + "\n")
    words.write_text("w," + " ".join(["0.05"] * 20) + "\n")
    base = [",".join(events_data.flow_row(rnd, ips)) for _ in range(4096)]
    p = tmp_path / "big.csv"
    with open(p, "w") as fh:
        fh.write(events_data.FLOW_HEADER + "\n")
        for rep in range(300):
            fh.write("\n".join(base) + "\n")
    out = []
    for threads in ("1", "7"):
        monkeypatch.setenv("ONI_IO_THREADS", threads)
        with FlowEvents(str(p), str(docs), str(words)) as ev:
            assert ev.info()["rows"] == 4096 * 300
            out.append({k: v.tolist() for k, v in ev.cuts().items()})
    assert out[0] == out[1]
    ref = eo.flow_post(events_data.FLOW_HEADER + "\n" + "\n".join(base) + "\n", docs.read_text(), words.read_text(), 0.0)
    assert out[0] == ref["cuts"]  # repeating the sample leaves its empirical CDF unchanged


def test_word_counts_feed_lda_pre(tmp_path):
    """<type>_pre_lda.scala's doc_wc.dat, then oni_lda_pre on it: the chain ml_ops.sh:44-67 runs."""
    import subprocess
    from oni_ml_b200 import LDA_BIN
    from oni_ml_b200.lda import lda_pre
    flow, docs, words = events_data.make_flow(str(tmp_path), n=3000, seed=9)
    ref = eo.flow_word_counts(open(flow).read())
    out = tmp_path / "doc_wc.dat"
    with FlowEvents(flow) as ev:
        assert ev.word_counts(str(out)) == len(ref)
        with pytest.raises(_lib.OniError):
            ev.score_write(1.0, str(tmp_path / "x"))  # no result tables were given
    assert out.read_text().split("\n")[:-1] == ref
    assert sum(int(ln.rsplit(",", 1)[1]) for ln in ref) == 2 * len(eo.flow_post(open(flow).read(), "", "", 0.0)["rows"])
    lda_pre(str(tmp_path) + "/")
    model = (tmp_path / "model.dat").read_text().split("\n")[:-1]
    assert len(model) == len({ln.split(",")[0] for ln in ref})  # one document per IP
    assert sum(int(ln.split(" ")[0]) for ln in model) == len(ref)  # one pair per line, duplicates kept
    # dns, through the command line (host-only stage: runs without a GPU)
    dns, top, _, _ = events_data.make_dns(str(tmp_path), n=3000, seed=10)
    r = subprocess.run([LDA_BIN, "dns_pre", dns, top, str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert out.read_text().split("\n")[:-1] == eo.dns_word_counts(open(dns).read(), open(top).read())


def test_analyst_feedback_rows_join_the_word_counts(tmp_path):
    """flow_scores.csv / dns_scores.csv rows of severity 3 come back as event rows, DUPFACTOR times
    each, behind the events of the day (flow_pre_lda.scala:147-263, dns_pre_lda.scala:70-138)."""
    rnd = random.Random(17)
    flow, _, _ = events_data.make_flow(str(tmp_path), n=1500, seed=3)
    fb = tmp_path / "flow_scores.csv"
    lines = ["sev,tstart,srcIP,dstIP,sport,dport,proto,flag,ipkt,ibyt,lda_score,rank,a,b,c,d,e,f,g,h,i,j"]
    for i in range(60):
        sev = rnd.choice([0, 1, 2, 3, 3])
        f = [str(sev), "2016-04-21 %02d:%02d:%02d" % (rnd.randint(0, 23), rnd.randint(0, 59), rnd.randint(0, 59)),
             "10.9.%d.1" % i, "10.8.%d.2" % i, str(rnd.choice([80, 443, 0, 50000])), str(rnd.choice([22, 0, 60000])),
             "TCP", ".A..", str(rnd.randint(1, 50)), str(rnd.randint(40, 90000)), "1e-9", str(i)] + ["x"] * 10
        if i % 13 == 0:
            f = f[:21]  # wrong column count: skipped
        lines.append(",".join(f))
    fb.write_text("\n".join(lines) + "\n")
    out = tmp_path / "doc_wc.dat"
    for dup in (0, 1, 3):
        rows = eo.flow_feedback_rows(fb.read_text(), dup)
        ref = eo.flow_word_counts(eo.with_rows(open(flow).read(), rows))
        with FlowEvents(flow, feedback_csv=str(fb), dup_factor=dup) as ev:
            assert ev.word_counts(str(out)) == len(ref)
            assert ev.info()["rows"] == len(eo.flow_post(open(flow).read(), "", "", 0.0)["rows"]) + len(rows)
        assert out.read_text().split("\n")[:-1] == ref
    assert len(eo.flow_feedback_rows(fb.read_text(), 1)) > 10
    bad = tmp_path / "bad.csv"
    bad.write_text(lines[0] + "\n" + ",".join(["three"] + ["x"] * 21) + "\n")
    with pytest.raises(_lib.OniError):
        FlowEvents(flow, feedback_csv=str(bad), dup_factor=1)
    # dns, through the command line
    import subprocess
    from oni_ml_b200 import LDA_BIN
    dns, top, _, _ = events_data.make_dns(str(tmp_path), n=1500, seed=4)
    fb2 = tmp_path / "dns_scores.csv"
    hdr = ",".join("c%d" % i for i in range(24))
    dl = [hdr]
    for i in range(40):
        f = ["Sep 13 2015 19:08:0%d UTC" % (i % 10), " %d " % rnd.choice([60, 74, 300]), "172.16.9.%d" % i,
             rnd.choice(events_data.DNS_NAMES[:6]), "0x00000001", "1", "0"] + ["y"] * 11 + [" %d" % rnd.choice([1, 3, 3]), "", "", "", "",
             str(1442171291 + i)]
        dl.append(",".join(f))
    fb2.write_text("\n".join(dl) + "\n")
    r = subprocess.run([LDA_BIN, "dns_pre", dns, top, str(out), str(fb2), "2"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    rows = eo.dns_feedback_rows(fb2.read_text(), 2)
    assert len(rows) > 20
    assert out.read_text().split("\n")[:-1] == eo.dns_word_counts(eo.with_rows(open(dns).read(), rows), open(top).read())


def test_dns_parquet_front_end(tmp_path):
    """DNS_PATH as the scripts take it: parquet files, null filter, eight columns, and the scripts'
    own habit of skipping the second file of the list (dns_pre_lda.scala:143-147)."""
    pa = pytest.importorskip("pyarrow")
    import pyarrow.parquet as pq
    from oni_ml_b200.events import DNS_COLUMNS, dns_parquet_to_csv
    paths = []
    for k in range(3):
        t = pa.table({
            "frame_time": ["t%d_%d" % (k, i) for i in range(4)], "unix_tstamp": [1442171291 + i if i != 1 else None for i in range(4)],
            "frame_len": [60 + i if i != 2 else None for i in range(4)], "ip_dst": ["10.0.%d.%d" % (k, i) for i in range(4)],
            "dns_qry_name": ["a.b.com", None, "c.org", "d.net"], "dns_qry_class": ["0x1"] * 4, "dns_qry_type": [1, 28, 1, 15],
            "dns_qry_rcode": [0, 0, 3, 0], "unused": [1.5] * 4})
        p = str(tmp_path / ("f%d.parquet" % k))
        pq.write_table(t, p)
        paths.append(p)
    out = str(tmp_path / "dns.csv")
    assert dns_parquet_to_csv(",".join(paths), out) == 4  # files 0 and 2, two rows each survive the null filter
    lines = open(out).read().split("\n")[:-1]
    assert lines[0] == "t0_0,1442171291,60,10.0.0.0,a.b.com,0x1,1,0" and lines[1].startswith("t0_3,1442171294,63,")
    assert all(ln.startswith(("t0_", "t2_")) for ln in lines) and len(lines[0].split(",")) == len(DNS_COLUMNS)
    assert dns_parquet_to_csv(",".join(paths), out, as_reference=False) == 6
