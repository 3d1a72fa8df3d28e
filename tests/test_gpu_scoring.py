"""Event scoring + ranking on the GPU against the oracle restatement
(oracle/score_oracle.c).  Integer and ordering results must be IDENTICAL; the
scores are compared bit for bit (unfused multiply/add in topic order on both
sides)."""
import numpy as np
import pytest

from oni_ml_b200 import scoring
from oracle import pyoracle

pytestmark = pytest.mark.gpu


def tables(D, V, K, seed):
    rng = np.random.default_rng(seed)
    theta = rng.dirichlet(np.full(K, 0.3), size=D)
    pwz = rng.dirichlet(np.full(V, 0.05), size=K).T.copy()
    return np.ascontiguousarray(theta), np.ascontiguousarray(pwz)


def events(n, D, V, seed, miss=0.01):
    rng = np.random.default_rng(seed)
    out = []
    for lim in (D, D, V, V):
        a = rng.integers(0, lim, size=n).astype(np.int32)
        a[rng.random(n) < miss] = -1
        out.append(a)
    return out  # src_doc, dst_doc, src_word, dst_word


def oracle_flow(theta, pwz, sd, dd, sw, dw, tol):
    L = pyoracle.lib()
    n, K = sd.shape[0], theta.shape[1]
    a = np.zeros(n); b = np.zeros(n)
    L.orc_score_flow(n, K, theta, pwz, sd, dd, sw, dw, a, b)
    idx = np.zeros(max(n, 1), np.int64); key = np.zeros(max(n, 1))
    m = L.orc_rank_below(n, a, b.ctypes.data, tol, n, idx, key)
    return a, b, idx[:m], key[:m]


@pytest.mark.parametrize("n,D,V,K,q", [(1000, 50, 40, 20, 0.3), (200000, 5000, 3000, 20, 0.01),
                                       (50000, 300, 200, 7, 1.0), (4097, 10, 10, 50, 0.5)])
def test_flow_ranked_list_identical(n, D, V, K, q):
    theta, pwz = tables(D, V, K, 1)
    sd, dd, sw, dw = events(n, D, V, 2)
    a, b, _, _ = oracle_flow(theta, pwz, sd, dd, sw, dw, np.inf)
    tol = float(np.quantile(np.minimum(a, b), q)) if q < 1.0 else 10.0
    _, _, want_idx, want_key = oracle_flow(theta, pwz, sd, dd, sw, dw, tol)
    got = scoring.score_flow(theta, pwz, sd, dd, sw, dw, tol)
    assert got["n"] == want_idx.shape[0]
    assert np.array_equal(got["idx"], want_idx)
    assert np.array_equal(got["key"].view(np.int64), want_key.view(np.int64))
    assert np.array_equal(got["src"].view(np.int64), a[want_idx].view(np.int64))
    assert np.array_equal(got["dst"].view(np.int64), b[want_idx].view(np.int64))
    assert np.all(np.diff(got["key"]) >= 0)


@pytest.mark.parametrize("chunk", [977, 4096, 65536])
def test_chunked_upload_gives_the_same_list_and_sides(monkeypatch, chunk):
    """Events from host arrays travel in chunks on a copy stream while the previous chunk is scored
    (ONI_SCORE_CHUNK; 2^24 by default).  Any chunk size, also one that does not divide the event
    count, must give the oracle's list -- indices, keys and the two per-side scores of the ranked
    survivors (which are gathered on the host across chunks)."""
    monkeypatch.setenv("ONI_SCORE_CHUNK", str(chunk))
    n, D, V, K = 30011, 700, 500, 20
    theta, pwz = tables(D, V, K, 5)
    sd, dd, sw, dw = events(n, D, V, 6)
    a, b, _, _ = oracle_flow(theta, pwz, sd, dd, sw, dw, np.inf)
    tol = float(np.quantile(np.minimum(a, b), 0.05))
    _, _, want_idx, want_key = oracle_flow(theta, pwz, sd, dd, sw, dw, tol)
    got = scoring.score_flow(theta, pwz, sd, dd, sw, dw, tol)
    assert got["n"] == want_idx.shape[0] and np.array_equal(got["idx"], want_idx)
    assert np.array_equal(got["key"].view(np.int64), want_key.view(np.int64))
    assert np.array_equal(got["src"].view(np.int64), a[want_idx].view(np.int64))
    assert np.array_equal(got["dst"].view(np.int64), b[want_idx].view(np.int64))
    # the context entry point a repeat caller uses: same events twice on one context, second run from a clean list
    ctx = scoring.ScoreContext(theta, pwz, 0)
    for _ in range(2):
        ctx.begin(); ctx.score_flow_host(sd, dd, sw, dw, tol); m = ctx.rank()
        idx, key = ctx.fetch(m)
        assert m == want_idx.shape[0] and np.array_equal(idx, want_idx)
        assert np.array_equal(key.view(np.int64), want_key.view(np.int64))
    ctx.close()
    # an index beyond its table, in the LAST chunk: refused by the kernel's own check
    bad = sd.copy(); bad[-3] = D
    from oni_ml_b200 import _lib
    with pytest.raises(_lib.OniError) as e:
        scoring.score_flow(theta, pwz, bad, dd, sw, dw, tol)
    assert e.value.code == 1


def test_ties_rank_by_event_index_and_misses_use_default_vector():
    K = 20
    theta = np.full((3, K), 0.05); pwz = np.full((4, K), 0.05)
    pwz[1] = np.linspace(0.001, 0.02, K)
    n = 5000
    sd = np.zeros(n, np.int32); dd = np.full(n, -1, np.int32)
    sw = np.zeros(n, np.int32); dw = np.full(n, -1, np.int32)
    sw[::7] = 1
    got = scoring.score_flow(theta, pwz, sd, dd, sw, dw, 1.0)
    a, b, want_idx, want_key = oracle_flow(theta, pwz, sd, dd, sw, dw, 1.0)
    assert got["n"] == n and np.array_equal(got["idx"], want_idx)
    # all-miss side: 20 * (0.05*0.05) accumulated sequentially
    acc = 0.0
    for _ in range(K):
        acc = acc + 0.05 * 0.05
    assert np.all(got["dst"] == acc)
    # ties (identical keys) come back in ascending event order
    same = got["key"][1:] == got["key"][:-1]
    assert same.any() and np.all(np.diff(got["idx"])[same] > 0)


def test_threshold_is_strict_and_cap_truncates():
    theta, pwz = tables(20, 30, 20, 5)
    sd, dd, sw, dw = events(3000, 20, 30, 6, miss=0.0)
    a, b, idx, key = oracle_flow(theta, pwz, sd, dd, sw, dw, np.inf)
    tol = float(key[100])  # exactly the 101st smallest key: strict '<' keeps 100 (no ties expected)
    got = scoring.score_flow(theta, pwz, sd, dd, sw, dw, tol, out_cap=10)
    assert got["n"] == int(np.sum(np.minimum(a, b) < tol))
    assert np.array_equal(got["idx"], idx[:10])
    none = scoring.score_flow(theta, pwz, sd, dd, sw, dw, 1e-20)  # the contract threshold: nothing survives
    assert none["n"] == 0 and none["idx"].shape == (0,)
    nothing = scoring.score_flow(theta, pwz, sd[:0], dd[:0], sw[:0], dw[:0], 1.0)
    assert nothing["n"] == 0


def test_dns_ranked_list_identical():
    theta, pwz = tables(400, 900, 20, 9)
    rng = np.random.default_rng(10)
    n = 123457
    doc = rng.integers(-1, 400, size=n).astype(np.int32)
    word = rng.integers(-1, 900, size=n).astype(np.int32)
    L = pyoracle.lib()
    s = np.zeros(n)
    L.orc_score_dns(n, 20, theta, pwz, doc, word, s)
    tol = float(np.quantile(s, 0.05))
    idx = np.zeros(n, np.int64); key = np.zeros(n)
    m = L.orc_rank_below(n, s, None, tol, n, idx, key)
    got = scoring.score_dns(theta, pwz, doc, word, tol)
    assert got["n"] == m and np.array_equal(got["idx"], idx[:m])
    assert np.array_equal(got["key"].view(np.int64), key[:m].view(np.int64))


def test_bad_event_index_is_rejected():
    theta, pwz = tables(5, 5, 20, 1)
    from oni_ml_b200 import _lib
    with pytest.raises(_lib.OniError):
        scoring.score_dns(theta, pwz, [7], [0], 1.0)


def test_lda_post_tables_match_oracle(golden_dir):
    """theta / p(w|z) shaping of lda_post.py from the reference's final.* files."""
    import os
    from oni_ml_b200 import files
    g = os.path.join(golden_dir, "c0_k20")
    gamma = files.read_matrix(os.path.join(g, "final.gamma"))
    beta = files.read_matrix(os.path.join(g, "final.beta"))
    L = pyoracle.lib()
    th = np.zeros_like(gamma); pw = np.zeros((beta.shape[1], beta.shape[0]))
    L.orc_theta_from_gamma(gamma.shape[0], gamma.shape[1], gamma, th)
    L.orc_pwz_from_beta(beta.shape[0], beta.shape[1], beta, pw)
    theta, pwz = scoring.lda_post(gamma, beta, 12)
    assert np.array_equal(theta, th)
    # exp() on the device and in glibc may differ in the last bit before the
    # 12-digit rounding; allow one unit of the 12th digit
    assert np.max(np.abs(pwz - pw) / pw) < 2e-12
    assert np.allclose(theta.sum(1), 1.0, atol=1e-10)


def test_lda_post_arithmetic_matches_the_reference_script(golden_dir):
    """oni_lda_post with text_digits=0 against tests/golden/post: the values the reference's own
    lda_post.py lines (:45-50, :88-96) produce under Python 3 (oracle/pin_lda_post.py)."""
    import os
    from oni_ml_b200 import files
    g = os.path.join(golden_dir, "c0_k20")
    gamma = files.read_matrix(os.path.join(g, "final.gamma"))
    beta = files.read_matrix(os.path.join(g, "final.beta"))
    want_theta = np.load(os.path.join(golden_dir, "post", "theta.npy"))
    want_pwz = np.load(os.path.join(golden_dir, "post", "pwz.npy"))
    theta, pwz = scoring.lda_post(gamma, beta, 0)
    assert np.max(np.abs(theta - want_theta) / want_theta) < 1e-15
    assert np.max(np.abs(pwz - want_pwz) / want_pwz) < 2e-12  # the device exp is a few ulp from numpy's


def test_lda_post_files_csvs(golden_dir, tmp_path):
    """oni_lda_post_files: doc_results.csv / word_results.csv in lda_post.py's shapes, values
    equal to the oracle's 12-digit tables, names cut to 20 bytes, ".0" on integral values."""
    import os
    import shutil
    from oni_ml_b200 import files
    g = os.path.join(golden_dir, "c0_k20")
    for n in ("final.gamma", "final.beta"):
        shutil.copy(os.path.join(g, n), tmp_path / n)
    gamma = files.read_matrix(os.path.join(g, "final.gamma"))
    beta = files.read_matrix(os.path.join(g, "final.beta"))
    D, V = gamma.shape[0], beta.shape[1]
    with open(tmp_path / "doc.dat", "w") as f:
        for i in range(D):
            f.write("%d,10.0.%d.%d\n" % (i + 1, i // 256, i % 256))
    with open(tmp_path / "words.dat", "w") as f:
        for w in range(V):
            f.write("%d,%s\n" % (w, "65535.0_9.0_9.0_9.0_tail" if w == 3 else "a,b" if w == 4 else "%d.0_1.0_2.0_3.0" % w))
    # a zero gamma row exercises the script's literal fallback
    lines = open(tmp_path / "final.gamma").read().splitlines()
    lines[5] = " ".join(["0.0000000000"] * 20)
    open(tmp_path / "final.gamma", "w").write("\n".join(lines) + "\n")
    gamma[5] = 0.0
    scoring.lda_post_files(str(tmp_path) + "/")
    L = pyoracle.lib()
    th = np.zeros_like(gamma); pw = np.zeros((V, 20))
    L.orc_theta_from_gamma(D, 20, gamma, th)
    L.orc_pwz_from_beta(20, V, beta, pw)
    rows = open(tmp_path / "doc_results.csv").read().splitlines()
    assert len(rows) == D and rows[0].startswith("10.0.0.0,") and rows[5] == "10.0.0.5," + " ".join(["0.0"] * 20)
    got = np.array([[float(x) for x in r.split(",")[1].split(" ")] for r in rows])
    assert np.array_equal(got, th)
    wrows = open(tmp_path / "word_results.csv").read().splitlines()
    assert len(wrows) == V
    assert wrows[3].split(",")[0] == "65535.0_9.0_9.0_9.0_"[:20] and wrows[4].split(",")[0] == "a_b"
    gotw = np.array([[float(x) for x in r.split(",")[1].split(" ")] for r in wrows])
    assert np.max(np.abs(gotw - pw) / pw) < 2e-12
    assert all(("." in t or "e" in t) for r in rows[:50] for t in r.split(",")[1].split(" "))
