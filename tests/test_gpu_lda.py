"""Parity of the CUDA engine (through the C ABI) with the oracle and with the
reference's own outputs (tests/golden, written by the unmodified oni-lda-c).

Tolerance: north_star asks for 1e-6 relative in fp64 on final.beta, final.gamma,
the likelihood trajectory and alpha; the engine evaluates the same recurrences
with a different (mathematically equal) association, so the observed agreement
is far tighter and the tests assert REL = 1e-6 on the golden comparisons and a
stricter 1e-9 against the oracle on single E-steps.  Index mappings, the
initial model (000.beta) and -100 floor entries are compared exactly.
"""
import os

import numpy as np
import pytest

from oni_ml_b200 import _lib, files, lda
from harness.synth import synth_corpus
from oni_ml_b200.corpus import Corpus
from oracle import pyoracle

pytestmark = pytest.mark.gpu
REL = 1e-6


def rel_err(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300))) if a.size else 0.0


def golden_case(golden_dir, case):
    d = os.path.join(golden_dir, case)
    args = dict(line.split() for line in open(os.path.join(d, "args.txt")))
    st = lda.read_settings(os.path.join(d, "settings.txt"))
    return d, args, st


def test_initial_model_byte_identical(golden_dir, c0_corpus, tmp_path):
    """MT19937 stream + random_initialize_ss + lda_mle: 000.beta equals the reference's bytes."""
    d, args, st = golden_case(golden_dir, "c0_k20")
    with lda.LdaEngine(0, 20, c0_corpus.num_terms, st) as eng:
        eng.init_random(int(args["seed"]), 2.5)
        lb0 = eng.logbeta()
    lda.save_model(str(tmp_path / "000"), lb0, 2.5)
    assert open(tmp_path / "000.beta").read() == open(os.path.join(d, "000.beta")).read()
    assert open(tmp_path / "000.other").read() == open(os.path.join(d, "000.other")).read()


@pytest.mark.parametrize("k,alpha,vmax", [(20, 2.5, 20), (5, 0.1, 5), (37, 1.0, -1), (8, 0.5, 20),
                                          (50, 0.3, 20), (3, 2.0, 20), (64, 1.5, 10), (1, 1.0, 5),
                                          (2, 0.7, 20), (12, 0.9, 20), (100, 0.2, 8), (200, 0.05, 6),
                                          (256, 1.0, 4), (20, 0.004, 20)])
def test_single_e_step_matches_oracle(c0_corpus, k, alpha, vmax):
    """One E-step from the same model: gamma, per-document bound, and the M-step's
    beta (i.e. the reduced class_word / class_total) against the oracle."""
    # (with one topic the reference's alpha Newton is 0/0 -- its alpha_suffstats is exactly 0 --
    #  so that case runs with "alpha fixed")
    st = _lib.Settings(vmax, 1e-6, 20, 1e-6, 0 if k == 1 else 1)
    with lda.LdaEngine(0, k, c0_corpus.num_terms, st) as eng:
        eng.set_corpus(c0_corpus)
        eng.init_random(1234, alpha)
        lb0 = eng.logbeta()
        ref = pyoracle.e_step(c0_corpus, k, alpha, lb0, vmax, 1e-6)
        eng.e_step()
        lik, conv = eng.m_step()
        gam = eng.gamma()
        lb1 = eng.logbeta()
        a1 = eng.alpha()
        stats = eng.stats()
    assert rel_err(gam, ref["gamma"]) < 1e-9
    assert abs(lik - ref["lik"]) / abs(ref["lik"]) < 1e-11
    assert np.isinf(conv) and conv > 0  # (0 - L)/0, lda-estimate.c:260
    L = pyoracle.lib()
    want = np.zeros_like(lb0)
    L.orc_mle_beta(k, c0_corpus.num_terms, ref["class_word"], ref["class_total"], want)
    assert np.array_equal(lb1 == -100.0, want == -100.0)
    assert rel_err(lb1, want) < 1e-9
    if k > 1:
        assert abs(a1 - L.orc_opt_alpha(ref["alpha_ss"], ref["num_docs"], k)) / a1 < 1e-9
    else:
        assert a1 == alpha
    assert stats["word_visits"] == ref["word_visits"]


@pytest.mark.parametrize("k,alpha", [(20, 2.5), (50, 0.3), (100, 0.2), (20, 0.004), (80, 0.4), (12, 0.9), (32, 1.1)])
def test_long_document_paths_match_oracle(monkeypatch, k, alpha):
    """The one-document-per-warp build (rows through the cp.async ring, requested 8/4/2 visits ahead
    for 1/2/4 topics per lane) and its exclusive-SM launch, forced onto documents of 64..700 pairs
    (lengths right at the ring's wrap-around limits included), beside ordinary short documents."""
    monkeypatch.setenv("ONI_LDA_LONG_DOC", "64")
    monkeypatch.setenv("ONI_LDA_EXCL_DOC", "300")
    rng = np.random.default_rng(11)
    V = 400
    lens = [64, 65, 66, 71, 72, 80, 127, 128, 129, 300, 301, 450, 700] + rng.integers(1, 64, 120).tolist()
    ptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    word = rng.integers(0, V, int(ptr[-1])).astype(np.int32)
    word[:V] = np.arange(V, dtype=np.int32) % V  # every word id occurs
    count = rng.geometric(0.3, int(ptr[-1])).astype(np.int32)
    c = Corpus(ptr, word, count, V)
    st = _lib.Settings(20, 1e-6, 20, 1e-6, 1)
    with lda.LdaEngine(0, k, V, st) as eng:
        eng.set_corpus(c)
        eng.init_random(99, alpha)
        lb0 = eng.logbeta()
        ref = pyoracle.e_step(c, k, alpha, lb0, 20, 1e-6)
        lik, _ = eng.em_iteration()
        gam, it1 = eng.gamma(), eng.doc_iters()
        lik2, _ = eng.em_iteration()  # a second E-step on the same handle: rings and tickets re-primed
        launches = eng.stats()["kernel_launches"]
    assert rel_err(gam, ref["gamma"]) < 1e-9
    assert abs(lik - ref["lik"]) / abs(ref["lik"]) < 1e-11
    assert int((np.diff(ptr) * it1).sum()) == ref["word_visits"]
    assert lik2 > lik and launches > 0


def test_long_document_leaves_the_digamma_table(monkeypatch):
    """The long-document kernel tabulates the reference's digamma on [2^-8, 2^23); a document
    whose gamma leaves that range (here: counts of ~10^6 on 200 pairs, gamma up to ~10^8) is run
    again with the series itself.  Both long documents (one inside the table, one outside) and
    the short ones beside them must match the oracle."""
    monkeypatch.setenv("ONI_LDA_LONG_DOC", "64")
    monkeypatch.setenv("ONI_LDA_EXCL_DOC", "150")
    rng = np.random.default_rng(5)
    V, k, alpha = 300, 20, 0.7
    lens = [200, 120, 90] + rng.integers(1, 40, 60).tolist()
    ptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    word = rng.integers(0, V, int(ptr[-1])).astype(np.int32)
    word[200:200 + V] = np.arange(V, dtype=np.int32)[: int(ptr[-1]) - 200][:V] % V
    count = rng.geometric(0.3, int(ptr[-1])).astype(np.int32)
    count[:200] = rng.integers(500_000, 2_000_000, 200).astype(np.int32)  # the first document
    c = Corpus(ptr, word, count, V)
    st = _lib.Settings(20, 1e-6, 20, 1e-6, 1)
    with lda.LdaEngine(0, k, V, st) as eng:
        eng.set_corpus(c)
        eng.init_random(7, alpha)
        lb0 = eng.logbeta()
        ref = pyoracle.e_step(c, k, alpha, lb0, 20, 1e-6)
        lik, _ = eng.em_iteration()
        gam, it1 = eng.gamma(), eng.doc_iters()
        s = eng.stats()
    assert s["docs_exclusive"] == 1 and s["docs_long"] == 2
    assert gam[0].max() > 2.0 ** 23  # the case the test is about
    assert rel_err(gam, ref["gamma"]) < 1e-9
    assert abs(lik - ref["lik"]) / abs(ref["lik"]) < 1e-11
    assert int((np.diff(ptr) * it1).sum()) == ref["word_visits"]


def test_tier_thresholds_follow_the_visit_count(monkeypatch):
    """After the first EM iteration the engine re-derives which documents run on the long-document
    launches from the shard's visit count (retier_from_visits, csrc/engine.cu).  That is scheduling
    only: the run with it and the run with the pair-count rule alone (ONI_LDA_RETIER=0) must give the
    same bound, alpha and gamma, although some documents change kernels between iterations."""
    monkeypatch.delenv("ONI_LDA_LONG_DOC", raising=False)
    monkeypatch.delenv("ONI_LDA_EXCL_DOC", raising=False)
    V, k = 20000, 20
    c = synth_corpus(200_000, V, mu=3.5, seed=77)
    st = _lib.Settings(20, 1e-6, 20, 1e-6, 1)
    runs = {}
    for mode in ("0", "1"):
        monkeypatch.setenv("ONI_LDA_RETIER", mode)
        with lda.LdaEngine(0, k, V, st) as eng:
            eng.set_corpus(c)
            eng.init_random(3, 2.5)
            liks, tiers = [], []
            for _ in range(3):
                liks.append(eng.em_iteration()[0])
                s = eng.stats()
                tiers.append((s["docs_exclusive"], s["docs_long"]))
            runs[mode] = (liks, tiers, eng.gamma(), eng.alpha(), s["word_visits"])
    (l0, t0, g0, a0, v0), (l1, t1, g1, a1, v1) = runs["0"], runs["1"]
    assert t0[0] == t0[1] == t0[2] and sum(t0[0]) > 0  # the pair-count rule does not move
    assert t1[0] == t0[0]                               # first iteration: nothing known yet
    assert t1[1] != t0[1]                               # then the visit count decides
    assert v0 == v1
    assert rel_err(l1, l0) < 1e-11 and abs(a1 - a0) / a0 < 1e-10
    assert rel_err(g1, g0) < 1e-8


def test_loaded_model_with_an_underflowing_word(c0_corpus):
    """A model that did not come from this engine's M-step (set_model / `lda inf` / a model-root
    start) may hold a word whose every topic sits far below the -100 floor.  exp(digamma + log
    beta) then underflows for all topics at once; the reference's log_sum chain
    (lda-inference.c:53-64) still normalises it, and so must the engine (it picks the
    max-shifted kernel from the model, not from alpha alone)."""
    k, alpha = 20, 2.5
    st = _lib.Settings(20, 1e-6, 20, 1e-6, 1)
    rng = np.random.default_rng(3)
    with lda.LdaEngine(0, k, c0_corpus.num_terms, st) as eng:
        eng.set_corpus(c0_corpus)
        eng.init_random(42, alpha)
        lb0 = eng.logbeta()
        used = np.unique(c0_corpus.word)
        for w in used[:: max(1, len(used) // 7)]:
            lb0[:, w] = -700.0 - 40.0 * rng.random(k)   # every topic of these words underflows
        eng.set_model(lb0, alpha)
        ref = pyoracle.e_step(c0_corpus, k, alpha, lb0, 20, 1e-6)
        eng.e_step()
        lik, _ = eng.m_step()
        gam = eng.gamma()
    assert np.isfinite(lik)
    assert rel_err(gam, ref["gamma"]) < 1e-9
    assert abs(lik - ref["lik"]) / abs(ref["lik"]) < 1e-11


def _check_against_golden(d, r, k, v, names):
    if "likelihood.dat" in names:
        want = files.read_likelihood(os.path.join(d, "likelihood.dat"))
        assert len(r["lik"]) == want.shape[0]
        assert rel_err(r["lik"], want[:, 0]) < REL
        assert np.isinf(r["conv"][0])
        # the printed relative change has 6 significant digits
        assert np.allclose(r["conv"][1:], want[1:, 1], rtol=2e-5, atol=1e-12)
    if "final.other" in names:
        o = files.read_other(os.path.join(d, "final.other"))
        assert (o["num_topics"], o["num_terms"]) == (k, v)
        assert abs(r["alpha"] - o["alpha"]) <= REL * o["alpha"] + 1e-10
    if "final.gamma" in names:
        want = files.read_matrix(os.path.join(d, "final.gamma"))
        assert np.max(np.abs(r["gamma"] - want) / (np.abs(want) + 1e-4)) < REL
    if "final.beta" in names:
        want = files.read_matrix(os.path.join(d, "final.beta"))
        assert np.array_equal(np.round(r["logbeta"], 10) == -100.0, want == -100.0)
        assert np.max(np.abs(r["logbeta"] - want) / (np.abs(want) + 1e-4)) < REL


def test_full_em_matches_reference_c0_k20(golden_dir, c0_corpus):
    d, args, st = golden_case(golden_dir, "c0_k20")
    r = lda.estimate(c0_corpus, 20, float(np.float32(2.5)), st, int(args["seed"]))
    assert r["iters"] == 21
    _check_against_golden(d, r, 20, c0_corpus.num_terms,
                          ["likelihood.dat", "final.other", "final.gamma", "final.beta"])
    want0 = files.read_matrix(os.path.join(d, "000.beta"))
    assert np.array_equal(np.round(r["logbeta0"], 10), want0)


def test_full_em_matches_three_rank_reference(golden_dir, c0_corpus):
    """The 3-rank reference run differs from ours only in summation order."""
    d, args, st = golden_case(golden_dir, "c0_k20_np3")
    r = lda.estimate(c0_corpus, 20, float(np.float32(2.5)), st, int(args["seed"]))
    _check_against_golden(d, r, 20, c0_corpus.num_terms, ["likelihood.dat", "final.other"])


def test_full_em_fixed_alpha_k5(golden_dir, c0_corpus):
    d, args, st = golden_case(golden_dir, "c0_k5_fixed")
    r = lda.estimate(c0_corpus, 5, float(np.float32(0.1)), st, int(args["seed"]))
    assert r["alpha"] == float(np.float32(0.1))
    _check_against_golden(d, r, 5, c0_corpus.num_terms, ["likelihood.dat", "final.other", "final.gamma"])


def test_full_em_unbounded_k37(golden_dir, c0_corpus):
    d, args, st = golden_case(golden_dir, "c0_k37_unbounded")
    r = lda.estimate(c0_corpus, 37, 1.0, st, int(args["seed"]))
    _check_against_golden(d, r, 37, c0_corpus.num_terms, ["likelihood.dat", "final.other"])


def test_cli_drop_in_writes_reference_files(golden_dir, tmp_path):
    """`lda est ...` with the reference's argv: same files, same formats."""
    d, args, st = golden_case(golden_dir, "c0_k20")
    out = tmp_path / "day"
    p = lda.run_cli(["est", "2.5", "20", os.path.join(d, "settings.txt"), "1",
                     os.path.join(golden_dir, "c0_model.dat"), "random", str(out)], seed=int(args["seed"]))
    assert "**** em iteration 21 ****" in p.stdout
    for name in ("000.beta", "000.other"):
        assert open(out / name).read() == open(os.path.join(d, name)).read(), name
    lik = files.read_likelihood(str(out / "likelihood.dat"))
    r = dict(lik=lik[:, 0], conv=lik[:, 1], alpha=files.read_other(str(out / "final.other"))["alpha"],
             gamma=files.read_matrix(str(out / "final.gamma")),
             logbeta=files.read_matrix(str(out / "final.beta")))
    _check_against_golden(d, r, 20, 300, ["likelihood.dat", "final.other", "final.gamma", "final.beta"])
    # text shape: K lines of V values with a leading blank; D lines of K values
    first = open(out / "final.beta").readline()
    assert first.startswith(" -") and len(first.split()) == 300
    assert len(open(out / "final.gamma").readlines()) == 403


def test_cli_inference_matches_oracle(golden_dir, c0_corpus, tmp_path):
    """`lda inf`: load_lda_model reads beta and alpha through float (lda-model.c:225-245)."""
    g = os.path.join(golden_dir, "c0_k20")
    stp = tmp_path / "inf-settings.txt"
    stp.write_text("var max iter -1\nvar convergence 1e-6\nem max iter 100\nem convergence 1e-4\nalpha estimate\n")
    lda.run_cli(["inf", str(stp), os.path.join(g, "final"), os.path.join(golden_dir, "c0_model.dat"),
                 str(tmp_path / "c0")])
    beta = files.read_matrix(os.path.join(g, "final.beta")).astype(np.float32).astype(np.float64)
    alpha = float(np.float32(files.read_other(os.path.join(g, "final.other"))["alpha"]))
    ref = pyoracle.e_step(c0_corpus, 20, alpha, beta, -1, 1e-6)
    gam = files.read_matrix(str(tmp_path / "c0-gamma.dat"))
    assert np.max(np.abs(gam - ref["gamma"]) / (np.abs(ref["gamma"]) + 1e-4)) < REL
    lh = np.loadtxt(str(tmp_path / "c0-lda-lhood.dat"))
    assert lh.shape == (403,)


def test_runs_are_bit_reproducible(c0_corpus):
    """Deterministic reductions: two runs give identical bits (no float atomics)."""
    st = _lib.Settings(20, 1e-6, 4, 1e-6, 1)
    a = lda.estimate(c0_corpus, 20, 2.5, st, 7)
    b = lda.estimate(c0_corpus, 20, 2.5, st, 7)
    for key in ("logbeta", "gamma", "lik"):
        assert np.array_equal(a[key], b[key]), key
    assert a["alpha"] == b["alpha"]


def test_config1_shape_properties_and_oracle_first_iterations():
    """BASELINE config 1 shape (10k docs, 5k words, K=20): first two EM iterations
    against the oracle, then size-independent invariants of the E-step."""
    c = synth_corpus(10000, 5000, mu=2.5, seed=20160123)
    st = _lib.Settings(20, 1e-6, 20, 1e-6, 1)
    with lda.LdaEngine(0, 20, 5000, st) as eng:
        eng.set_corpus(c)
        eng.init_random(4357, 2.5)
        lb0 = eng.logbeta()
        lik1, _ = eng.em_iteration()
        gam1, a1 = eng.gamma(), eng.alpha()
        lb1 = eng.logbeta()
        lik2, conv2 = eng.em_iteration()
        iters = eng.doc_iters()
    ref = pyoracle.e_step(c, 20, 2.5, lb0, 20, 1e-6)
    assert abs(lik1 - ref["lik"]) / abs(ref["lik"]) < 1e-10
    assert rel_err(gam1, ref["gamma"]) < 1e-8
    ref2 = pyoracle.e_step(c, 20, a1, lb1, 20, 1e-6)
    assert abs(lik2 - ref2["lik"]) / abs(ref2["lik"]) < 1e-10
    assert 0 < conv2 < 1 and lik2 > lik1
    # invariants: sum_k gamma_dk = K*alpha + total_d ; exp(beta) rows sum to 1
    tot = np.add.reduceat(c.count.astype(np.float64), c.doc_ptr[:-1])
    assert np.allclose(gam1.sum(1), 20 * 2.5 + tot, rtol=1e-12)
    assert np.allclose(np.exp(lb1).sum(1), 1.0, rtol=1e-9)
    assert iters.min() >= 1 and iters.max() <= 20


def test_empty_and_degenerate_shards():
    """A shard with no documents, and documents with no words, are legal."""
    st = _lib.Settings(20, 1e-6, 3, 1e-6, 0)
    empty = Corpus(np.zeros(1, np.int64), np.zeros(0, np.int32), np.zeros(0, np.int32), 50)
    with lda.LdaEngine(0, 20, 50, st) as eng:
        eng.set_corpus(empty)
        eng.init_random(1, 1.0)
        eng.e_step()
        lik, _ = eng.m_step()
        assert lik == 0.0 and eng.gamma().shape == (0, 20)
        assert np.all(eng.logbeta() == -100.0)  # class_word == 0 everywhere (lda-model.c:42)
    blanks = Corpus(np.zeros(4, np.int64), np.zeros(0, np.int32), np.zeros(0, np.int32), 50)
    with lda.LdaEngine(0, 20, 50, st) as eng:
        eng.set_corpus(blanks)
        eng.init_random(1, 1.0)
        eng.e_step()
        eng.m_step()
        assert np.allclose(eng.gamma(), 1.0)  # alpha + 0/K


def test_bad_arguments_are_rejected():
    st = _lib.Settings.shipped()
    with pytest.raises(_lib.OniError):
        lda.LdaEngine(0, 0, 10, st)
    with pytest.raises(_lib.OniError):
        lda.LdaEngine(0, 300, 10, st)  # K > 256
    with lda.LdaEngine(0, 4, 10, st) as eng:
        with pytest.raises(_lib.OniError):
            eng.e_step()  # no model / corpus yet
        bad = Corpus(np.array([0, 1], np.int64), np.array([10], np.int32), np.array([1], np.int32), 10)
        with pytest.raises(_lib.OniError):
            eng.set_corpus(bad)  # word id out of range


def test_flow_day_size_invariants():
    """BASELINE configs[1] at full size (1M documents, 50k words, ~55M pairs, K=20): the
    oracle cannot run this in seconds, so the check is through size-independent
    properties of two EM iterations: sum_k gamma_dk = K*alpha + total_d, exp(beta)
    rows sum to 1, the bound increases, per-document bounds add up to the reduced
    total, the reductions are bit-reproducible across runs, and a 20k-document
    sample of documents matches the oracle's per-document inference."""
    import torch
    ptr, word, count = synth_corpus(1_000_000, 50_000, mu=3.5, seed=20160123, device="cuda", as_torch=True)
    st = _lib.Settings(20, 1e-6, 20, 1e-6, 1)
    runs = []
    for _ in range(2):
        with lda.LdaEngine(0, 20, 50_000, st) as eng:
            eng.set_corpus_device(ptr, word, count)
            eng.init_random(4357, 2.5)
            lb0 = eng.logbeta() if not runs else None
            lik1, _ = eng.em_iteration()
            gam1 = eng.gamma()
            doc_lik = eng.doc_likelihood()
            lb1, a1 = eng.logbeta(), eng.alpha()
            lik2, conv2 = eng.em_iteration()
            runs.append((lik1, lik2, gam1, lb1, a1))
            if lb0 is not None:
                keep = (lb0, doc_lik)
    (lik1, lik2, gam1, lb1, a1), other = runs
    assert lik1 == other[0] and lik2 == other[1] and a1 == other[4]
    assert np.array_equal(gam1, other[2]) and np.array_equal(lb1, other[3])
    assert lik2 > lik1 and 0 < conv2 < 1
    lb0, doc_lik = keep
    assert abs(doc_lik.sum() - lik1) < 1e-9 * abs(lik1)
    h_ptr, h_count = ptr.cpu().numpy(), count.cpu().numpy()
    tot = np.add.reduceat(h_count.astype(np.float64), h_ptr[:-1])
    assert np.allclose(gam1.sum(1), 20 * 2.5 + tot, rtol=1e-12)
    assert np.allclose(np.exp(lb1).sum(1), 1.0, rtol=1e-9)
    # oracle on a sample of documents (including the longest ones)
    lens = np.diff(h_ptr)
    pick = np.unique(np.concatenate([np.argsort(lens)[-20:], np.arange(0, 1_000_000, 50)]))
    h_word = word.cpu().numpy()
    L = pyoracle.lib()
    worst = 0.0
    for d in pick:
        a, b = int(h_ptr[d]), int(h_ptr[d + 1])
        g = np.zeros(20); phi = np.zeros((max(b - a, 1), 20))
        lk = L.orc_doc_inference(b - a, np.ascontiguousarray(h_word[a:b]), np.ascontiguousarray(h_count[a:b]),
                                 int(h_count[a:b].sum()), 20, 50_000, 2.5, lb0, 20, float(np.float32(1e-6)),
                                 g, phi, None)
        worst = max(worst, float(np.max(np.abs(gam1[d] - g) / g)), abs(doc_lik[d] - lk) / abs(lk))
    assert worst < 1e-8, worst


def test_config1_full_run_matches_reference_binary(golden_dir):
    """BASELINE configs[0] end to end: 10k documents, 5k words, K=20, alpha 2.5, shipped
    settings, 21 EM iterations -- against what the unmodified oni-lda-c wrote for the same
    corpus and seed (tests/golden/c1_k20, oracle/make_golden.py): the whole likelihood
    trajectory, alpha, every 50th row of final.gamma and every 25th column of final.beta."""
    d = os.path.join(golden_dir, "c1_k20")
    c = synth_corpus(10000, 5000, k_true=20, mu=2.5, seed=20160123)
    crc = open(os.path.join(d, "corpus.txt")).read().splitlines()[1].split()
    if (int(c.word.astype(np.int64).sum()), int(c.count.astype(np.int64).sum()), int(c.doc_ptr.sum())) != (
            int(crc[2]), int(crc[4]), int(crc[6])):
        pytest.skip("the synthetic generator draws differently on this torch build")
    st = lda.read_settings(os.path.join(d, "settings.txt"))
    r = lda.estimate(c, 20, float(np.float32(2.5)), st, 4357)
    want = files.read_likelihood(os.path.join(d, "likelihood.dat"))
    assert r["iters"] == want.shape[0] == 21
    assert rel_err(r["lik"], want[:, 0]) < REL
    o = files.read_other(os.path.join(d, "final.other"))
    assert abs(r["alpha"] - o["alpha"]) < REL * o["alpha"] + 1e-10
    gam = np.loadtxt(os.path.join(d, "final.gamma.rows50"))
    assert np.max(np.abs(r["gamma"][::50] - gam) / (np.abs(gam) + 1e-4)) < REL
    beta = np.loadtxt(os.path.join(d, "final.beta.cols25"))
    got = r["logbeta"][:, ::25]
    assert np.array_equal(np.round(got, 10) == -100.0, beta == -100.0)
    assert np.max(np.abs(got - beta) / (np.abs(beta) + 1e-4)) < REL


def test_flow_day_full_run_matches_reference_binary(golden_dir):
    """BASELINE configs[1], the size the headline is quoted on: 1M documents, 50k words, ~55M
    pairs, K=20, `est 2.5 20 settings.txt ... random` with the shipped settings -- all 21 EM
    iterations on the GPU against what the UNMODIFIED oni-lda-c wrote for the same corpus and seed
    (tests/golden/c2_flow_day, oracle/make_golden.py c2: the reference binary on 7 forked ranks,
    ~1.5 h of CPU).  Compared at 1e-6: every line of likelihood.dat (iterations 2..21 run with the
    re-estimated alpha and whatever var-max-iter doubling the run went through), alpha, every
    5000th row of final.gamma and every 250th column of final.beta; the -100 floor entries of beta
    exactly (lda-estimate.c:186-275)."""
    d = os.path.join(golden_dir, "c2_flow_day")
    if not os.path.exists(os.path.join(d, "likelihood.dat")):
        pytest.skip("tests/golden/c2_flow_day is not there")
    c = synth_corpus(1_000_000, 50_000, k_true=20, mu=3.5, seed=20160123)
    crc = open(os.path.join(d, "corpus.txt")).read().splitlines()[1].split()
    if (int(c.word.astype(np.int64).sum()), int(c.count.astype(np.int64).sum()), int(c.doc_ptr.sum())) != (
            int(crc[2]), int(crc[4]), int(crc[6])):
        pytest.skip("the synthetic generator draws differently on this torch build")
    st = lda.read_settings(os.path.join(d, "settings.txt"))
    r = lda.estimate(c, 20, float(np.float32(2.5)), st, 4357)
    want = files.read_likelihood(os.path.join(d, "likelihood.dat"))
    assert r["iters"] == want.shape[0]
    assert rel_err(r["lik"], want[:, 0]) < REL
    assert np.allclose(r["conv"][1:], want[1:, 1], rtol=2e-5, atol=1e-12)
    o = files.read_other(os.path.join(d, "final.other"))
    assert abs(r["alpha"] - o["alpha"]) < REL * o["alpha"] + 1e-10
    gam = np.loadtxt(os.path.join(d, "final.gamma.rows5000"))
    assert np.max(np.abs(r["gamma"][::5000] - gam) / (np.abs(gam) + 1e-4)) < REL
    beta = np.loadtxt(os.path.join(d, "final.beta.cols250"))
    got = r["logbeta"][:, ::250]
    assert np.array_equal(np.round(got, 10) == -100.0, beta == -100.0)
    assert np.max(np.abs(got - beta) / (np.abs(beta) + 1e-4)) < REL


def _cli_est(golden_dir, tmp_path, case, start, seed):
    d = os.path.join(golden_dir, case)
    out = tmp_path / case
    lda.run_cli(["est", "2.5", "20", os.path.join(d, "settings.txt"), "1",
                 os.path.join(golden_dir, "c0_model.dat"), start, str(out)], seed=seed)
    return d, out


def test_cli_seeded_start_matches_reference(golden_dir, tmp_path):
    """`seeded`: corpus_initialize_ss (lda-model.c:120-144) draws its documents from the same
    MT19937 stream -> 000.beta byte-identical, trajectory within 1e-6."""
    d, out = _cli_est(golden_dir, tmp_path, "c0_k20_seeded", "seeded", 777)
    assert open(out / "000.beta").read() == open(os.path.join(d, "000.beta")).read()
    got, want = files.read_likelihood(str(out / "likelihood.dat")), files.read_likelihood(os.path.join(d, "likelihood.dat"))
    assert got.shape == want.shape and rel_err(got[:, 0], want[:, 0]) < REL
    a, b = files.read_other(str(out / "final.other")), files.read_other(os.path.join(d, "final.other"))
    assert abs(a["alpha"] - b["alpha"]) < REL * b["alpha"] + 1e-10


def test_cli_resume_from_model_root_matches_reference(golden_dir, tmp_path):
    """Start from <root>.beta/.other: beta and alpha are re-read through float
    (load_lda_model, lda-model.c:220-251) and alpha comes from the file, not argv."""
    d, out = _cli_est(golden_dir, tmp_path, "c0_k20_resume", os.path.join(golden_dir, "c0_k20", "final"), 1)
    assert open(out / "000.other").read() == open(os.path.join(d, "000.other")).read()
    got, want = files.read_likelihood(str(out / "likelihood.dat")), files.read_likelihood(os.path.join(d, "likelihood.dat"))
    assert got.shape == want.shape and rel_err(got[:, 0], want[:, 0]) < REL
    a, b = files.read_other(str(out / "final.other")), files.read_other(os.path.join(d, "final.other"))
    assert abs(a["alpha"] - b["alpha"]) < REL * b["alpha"] + 1e-10


def test_cli_inference_matches_reference_binary(golden_dir, tmp_path):
    """`lda inf` against the reference's own <name>-lda-lhood.dat / <name>-gamma.dat."""
    d = os.path.join(golden_dir, "c0_k20_inf")
    lda.run_cli(["inf", os.path.join(d, "settings.txt"), os.path.join(golden_dir, "c0_k20", "final"),
                 os.path.join(golden_dir, "c0_model.dat"), str(tmp_path / "c0")])
    gam, want = files.read_matrix(str(tmp_path / "c0-gamma.dat")), files.read_matrix(os.path.join(d, "c0-gamma.dat"))
    assert np.max(np.abs(gam - want) / (np.abs(want) + 1e-4)) < REL
    lh, wl = np.loadtxt(str(tmp_path / "c0-lda-lhood.dat")), np.loadtxt(os.path.join(d, "c0-lda-lhood.dat"))
    assert np.allclose(lh, wl, rtol=REL, atol=2e-5)  # printed with 5 decimals (lda-estimate.c:387)
