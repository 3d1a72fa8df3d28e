"""Raw events -> scored rows through the C ABI (oni_*_events_open + oni_events_score_write,
and the `lda flow_post` / `lda dns_post` commands) against the Python restatement of the two
Spark scripts: the scored files must match byte for byte."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import events_oracle as eo  # noqa: E402
from test_events_cpu import _dns_case, _flow_case  # noqa: E402

from oni_ml_b200 import LDA_BIN  # noqa: E402
from oni_ml_b200.events import DnsEvents, FlowEvents  # noqa: E402

pytestmark = pytest.mark.gpu


def _quantile(scores, q):
    s = sorted(scores)
    return s[int(q * (len(s) - 1))]


def test_flow_scored_file(tmp_path):
    flow, docs, words = _flow_case(tmp_path, 6000, 21)
    args = (open(flow).read(), open(docs).read(), open(words).read())
    everything = eo.flow_post(*args, 1.0)
    keys = [min(float(r.split(",")[-2]), float(r.split(",")[-1])) for r in everything["scored"]]
    assert len(keys) == len(everything["rows"])  # every score is below 1
    out = str(tmp_path / "scored.csv")
    with FlowEvents(flow, docs, words) as ev:
        for tol in (1.0, _quantile(keys, 0.05), _quantile(keys, 0.5), 1e-20):
            ref = eo.flow_post(*args, tol)
            n = ev.score_write(tol, out)
            got = open(out).read().split("\n")
            assert got[-1] == "" and got[:-1] == ref["scored"], tol
            assert n == len(ref["scored"])
    # the command line replacing `spark-shell -i flow_post_lda.scala`
    tol = _quantile(keys, 0.1)
    r = subprocess.run([LDA_BIN, "flow_post", flow, str(tmp_path) + "/", repr(tol), out], capture_output=True,
                       text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert open(out).read().split("\n")[:-1] == eo.flow_post(*args, tol)["scored"]


def test_dns_scored_file(tmp_path):
    dns, top, docs, words = _dns_case(tmp_path, 6000, 22)
    args = (open(dns).read(), open(top).read(), open(docs).read(), open(words).read())
    everything = eo.dns_post(*args, 10.0)
    keys = [float(r.split(",")[-1]) for r in everything["scored"]]
    out = str(tmp_path / "scored.csv")
    with DnsEvents(dns, top, docs, words) as ev:
        for tol in (10.0, _quantile(keys, 0.05), _quantile(keys, 0.5)):
            ref = eo.dns_post(*args, tol)
            n = ev.score_write(tol, out)
            assert open(out).read().split("\n")[:-1] == ref["scored"], tol
            assert n == len(ref["scored"])
    tol = _quantile(keys, 0.1)
    r = subprocess.run([LDA_BIN, "dns_post", dns, top, str(tmp_path) + "/", repr(tol), out], capture_output=True,
                       text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert open(out).read().split("\n")[:-1] == eo.dns_post(*args, tol)["scored"]


def test_whole_flow_pipeline_from_the_command_line(tmp_path):
    """ml_ops.sh:44-108 with every stage replaced by the `lda` executable: raw flows -> doc_wc.dat ->
    words/doc/model.dat -> EM -> doc_results/word_results -> scored rows."""
    import events_data
    flow, _, _ = events_data.make_flow(str(tmp_path), n=5000, seed=31)
    for name in ("doc_results.csv", "word_results.csv"):
        p = tmp_path / name
        if p.exists():
            p.unlink()
    d = str(tmp_path) + "/"
    (tmp_path / "settings.txt").write_text("var max iter 20\nvar convergence 1e-6\nem max iter 5\n"
                                           "em convergence 1e-4\nalpha estimate\n")
    env = dict(os.environ, ONI_LDA_SEED="4357")
    for cmd in (["flow_pre", flow, d + "doc_wc.dat"], ["pre", d],
                ["est", "2.5", "20", d + "settings.txt", "1", d + "model.dat", "random", str(tmp_path)],
                ["post", d], ["flow_post", flow, d, "1.0", d + "scored.csv"]):
        r = subprocess.run([LDA_BIN] + cmd, capture_output=True, text=True, env=env)
        assert r.returncode == 0, (cmd, r.stdout, r.stderr)
    got = open(d + "scored.csv").read().split("\n")[:-1]
    ref = eo.flow_post(open(flow).read(), open(d + "doc_results.csv").read(), open(d + "word_results.csv").read(), 1.0)
    assert got == ref["scored"] and len(got) == len(ref["rows"]) > 4500
    keys = [min(float(r.split(",")[-2]), float(r.split(",")[-1])) for r in got]
    assert keys == sorted(keys) and 0 < keys[0] < keys[-1] < 1
    # every IP of the day got a topic mix; words longer than 20 bytes miss, as in the reference chain
    hit = sum(a and b for a, b in ref["doc_hits"])
    assert hit == len(ref["rows"])
