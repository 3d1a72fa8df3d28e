"""The whole post-LDA chain, reference side against GPU side, on BASELINE configs[0].

north_star: "the set and order of events below the suspicion threshold must be identical".
The two chains, on the same corpus, seed and 10^6 synthetic flow events:

  reference   final.gamma / final.beta as the unmodified oni-lda-c wrote them
              (tests/golden/c1_k20/final_full.npz = the doubles their "%5.10f" text parses to)
              -> lda_post.py shaping (oracle/score_oracle.c; its arithmetic is pinned to the
                 script's own lines by tests/golden/post, the 12-digit text is the restated
                 convention)
              -> flow_post_lda.scala:227-245 scoring and ranking (oracle, ties by event index)
  GPU         oni_lda_estimate (21 EM iterations) -> final.* through the library's "%5.10f"
              writers and back -> oni_lda_post (12 digits) -> oni_score_flow

gamma and beta of the two engines agree to ~1e-9, so after the 10-decimal and 12-digit text
steps most table entries are the same doubles and the rest differ in their last printed digit;
an event's score then moves by ~1e-11 relative.  The ranked list can only change where two
DIFFERENT scores lie closer together than that, or where a score sits that close to the
threshold.  The thresholds are put in the middle of a gap between neighbouring reference scores
near the 1e-3 and 1e-2 quantiles (as an analyst's round number would be), and the test then
asserts the identical set and the identical order, and reports the closest pair it saw.
"""
import os

import numpy as np
import pytest

from oni_ml_b200 import files, lda, scoring
from harness.synth import synth_corpus
from oracle import pyoracle

pytestmark = pytest.mark.gpu


def test_ranked_list_of_the_gpu_chain_equals_the_reference_chain(golden_dir, tmp_path):
    d = os.path.join(golden_dir, "c1_k20")
    c = synth_corpus(10000, 5000, k_true=20, mu=2.5, seed=20160123)
    crc = open(os.path.join(d, "corpus.txt")).read().splitlines()[1].split()
    if (int(c.word.astype(np.int64).sum()), int(c.count.astype(np.int64).sum()), int(c.doc_ptr.sum())) != (
            int(crc[2]), int(crc[4]), int(crc[6])):
        pytest.skip("the synthetic generator draws differently on this torch build")
    ref = np.load(os.path.join(d, "final_full.npz"))
    ref_gamma, ref_beta = ref["gamma"], ref["beta"]
    K, D, V = 20, c.num_docs, c.num_terms
    assert ref_gamma.shape == (D, K) and ref_beta.shape == (K, V)

    # ---- GPU chain: EM, then the files' own text and back
    st = lda.read_settings(os.path.join(d, "settings.txt"))
    r = lda.estimate(c, K, float(np.float32(2.5)), st, 4357)
    lda.save_model(str(tmp_path / "final"), r["logbeta"], r["alpha"])
    lda.save_gamma(str(tmp_path / "final.gamma"), r["gamma"])
    gpu_gamma = files.read_matrix(str(tmp_path / "final.gamma"))
    gpu_beta = files.read_matrix(str(tmp_path / "final.beta"))
    same_gamma = float(np.mean(gpu_gamma == ref_gamma))
    same_beta = float(np.mean(gpu_beta == ref_beta))
    assert np.max(np.abs(gpu_gamma - ref_gamma) / (np.abs(ref_gamma) + 1e-4)) < 1e-6
    gpu_theta, gpu_pwz = scoring.lda_post(gpu_gamma, gpu_beta, 12)

    # ---- reference chain: lda_post.py shaping of the reference's files
    L = pyoracle.lib()
    ref_theta = np.zeros_like(ref_gamma); ref_pwz = np.zeros((V, K))
    L.orc_theta_from_gamma(D, K, np.ascontiguousarray(ref_gamma), ref_theta)
    L.orc_pwz_from_beta(K, V, np.ascontiguousarray(ref_beta), ref_pwz)

    # ---- 10^6 synthetic flow events: uniform documents, Zipf-like words, 0.1 % lookup misses
    n = 1_000_000
    rng = np.random.default_rng(20160127)
    ev = [rng.integers(0, D, n).astype(np.int32), rng.integers(0, D, n).astype(np.int32),
          (V * rng.random(n) ** 3).astype(np.int32), (V * rng.random(n) ** 3).astype(np.int32)]
    for a in ev:
        a[rng.random(n) < 1e-3] = -1
    sa = np.zeros(n); sb = np.zeros(n)
    L.orc_score_flow(n, K, ref_theta, ref_pwz, *ev, sa, sb)
    ref_score = np.minimum(sa, sb)
    srt = np.sort(ref_score)

    report = []
    for q in (1e-3, 1e-2):
        k = int(q * n)
        # the widest gap among the 200 neighbouring pairs around the quantile: the threshold goes in its middle
        w = srt[k - 100:k + 101]
        j = int(np.argmax(np.diff(w)))
        tol = 0.5 * (w[j] + w[j + 1])
        oi = np.zeros(n, np.int64); ok = np.zeros(n)
        m = L.orc_rank_below(n, sa, sb.ctypes.data, tol, n, oi, ok)
        got = scoring.score_flow(gpu_theta, gpu_pwz, *ev, tol)
        want_idx, got_idx = oi[:m], got["idx"]
        only_ref = np.setdiff1d(want_idx, got_idx).size
        only_gpu = np.setdiff1d(got_idx, want_idx).size
        first_diff = -1
        if got["n"] == m and not np.array_equal(want_idx, got_idx):
            first_diff = int(np.argmax(want_idx != got_idx))
        # how close the closest pair of DIFFERENT neighbouring scores is, relative
        ks = ok[:m]
        gaps = np.diff(ks)
        rel_gap = gaps[gaps > 0] / ks[1:][gaps > 0]
        report.append(dict(q=q, tol=tol, survivors=int(m), only_ref=only_ref, only_gpu=only_gpu,
                           first_order_difference=first_diff,
                           closest_distinct_pair=float(rel_gap.min()) if rel_gap.size else None,
                           max_rel_score_diff=float(np.max(np.abs(got["key"] - ks) / ks)) if got["n"] == m else None))
    print("chain-level ranked list:", dict(gamma_entries_identical=same_gamma, beta_entries_identical=same_beta), report)
    for rep in report:
        assert rep["only_ref"] == 0 and rep["only_gpu"] == 0, report      # the same set
        assert rep["first_order_difference"] == -1, report                 # in the same order
        assert rep["max_rel_score_diff"] < 1e-9, report
