"""Pin the oracle restatement (oracle/lda_oracle.c) to the reference's own
outputs: the golden files under tests/golden/ were written by the UNMODIFIED
oni-lda-c sources (oracle/make_golden.py).  Bit-exact: the restatement keeps the
reference's order of evaluation, so the formatted text must be byte-identical.
"""
import os

import numpy as np
import pytest

from oni_ml_b200 import files
from oracle import pyoracle


def _args(d):
    out = {}
    with open(os.path.join(d, "args.txt")) as f:
        for line in f:
            k, v = line.split()
            out[k] = v
    with open(os.path.join(d, "settings.txt")) as f:
        st = pyoracle.parse_settings(f.read())
    return out, st


def _run(golden_dir, c0_corpus, case):
    d = os.path.join(golden_dir, case)
    a, st = _args(d)
    r = pyoracle.run_em(c0_corpus, int(a["k"]), float(a["alpha"]), st, int(a["seed"]),
                        int(a["nproc"]))
    return d, r, int(a["k"])


def _text(path):
    with open(path) as f:
        return f.read()


def test_c0_k20_all_files_byte_identical(golden_dir, c0_corpus):
    d, r, k = _run(golden_dir, c0_corpus, "c0_k20")
    v = c0_corpus.num_terms
    assert files.format_beta(r["logbeta0"]) == _text(os.path.join(d, "000.beta"))
    assert files.format_other(k, v, float(np.float32(2.5))) == _text(os.path.join(d, "000.other"))
    assert files.format_likelihood(r["lik"], r["conv"]) == _text(os.path.join(d, "likelihood.dat"))
    assert files.format_beta(r["logbeta"]) == _text(os.path.join(d, "final.beta"))
    assert files.format_gamma(r["gamma"]) == _text(os.path.join(d, "final.gamma"))
    assert files.format_other(k, v, r["alpha"]) == _text(os.path.join(d, "final.other"))


def test_c0_three_ranks(golden_dir, c0_corpus):
    """nproc=3: strided partition (lda-estimate.c:201) and rank-ordered sums."""
    d, r, k = _run(golden_dir, c0_corpus, "c0_k20_np3")
    assert files.format_likelihood(r["lik"], r["conv"]) == _text(os.path.join(d, "likelihood.dat"))
    assert files.format_other(k, c0_corpus.num_terms, r["alpha"]) == _text(os.path.join(d, "final.other"))


def test_c0_k5_fixed_alpha(golden_dir, c0_corpus):
    """alpha fixed, alpha parsed through a float (0.1 -> 0.1000000015)."""
    d, r, k = _run(golden_dir, c0_corpus, "c0_k5_fixed")
    assert files.format_likelihood(r["lik"], r["conv"]) == _text(os.path.join(d, "likelihood.dat"))
    assert files.format_gamma(r["gamma"]) == _text(os.path.join(d, "final.gamma"))
    assert files.format_other(k, c0_corpus.num_terms, r["alpha"]) == _text(os.path.join(d, "final.other"))


def test_c0_k37_unbounded_var_iters(golden_dir, c0_corpus):
    """var max iter -1 (no cap), K not a multiple of anything convenient."""
    d, r, k = _run(golden_dir, c0_corpus, "c0_k37_unbounded")
    assert files.format_likelihood(r["lik"], r["conv"]) == _text(os.path.join(d, "likelihood.dat"))
    assert files.format_other(k, c0_corpus.num_terms, r["alpha"]) == _text(os.path.join(d, "final.other"))


def test_mt19937_known_answer():
    """cokus.c is MT19937 with the 1998 seeding (LCG 69069); with seed 4357 its
    stream must open like the published mt19937int.out of that release.  (The
    reference itself holds no vector; 000.beta above pins the same stream.)"""
    L = pyoracle.lib()
    L.orc_mt_seed(4357)
    got = [L.orc_mt_next() for _ in range(5)]
    assert got == [3510405877, 4290933890, 2191955339, 564929546, 152112058], got


@pytest.mark.skipif(not os.path.exists(pyoracle.REF_BIN), reason="oracle/_ref not built")
def test_reference_binary_reproduces_golden(tmp_path, golden_dir):
    """The committed fixtures are what the reference binary writes today."""
    import subprocess
    d = os.path.join(golden_dir, "c0_k5_fixed")
    out = tmp_path / "out"
    env = dict(os.environ, ONI_LDA_SEED="99", ONI_MPI_NP="1")
    subprocess.run([pyoracle.REF_BIN, "est", "0.1", "5", os.path.join(d, "settings.txt"), "1",
                    os.path.join(golden_dir, "c0_model.dat"), "random", str(out)],
                   check=True, env=env, stdout=subprocess.DEVNULL)
    for name in ("likelihood.dat", "final.gamma", "final.other"):
        assert _text(out / name) == _text(os.path.join(d, name))


def test_post_shaping_arithmetic_pinned_to_the_reference_script(golden_dir):
    """tests/golden/post holds theta and p(w|z) as the reference's OWN source lines compute them
    (lda_post.py:45-50 and :88-96 executed under Python 3 by oracle/pin_lda_post.py, at full double
    precision).  The C restatement must give the same values: theta exactly (one division per
    entry), p(w|z) to the last bits (numpy's exp against glibc's).  The 12-digit text layer on top
    of these values stays a restated convention (no Python 2 here)."""
    from oni_ml_b200 import files
    g = os.path.join(golden_dir, "c0_k20")
    gamma = files.read_matrix(os.path.join(g, "final.gamma"))
    beta = files.read_matrix(os.path.join(g, "final.beta"))
    want_theta = np.load(os.path.join(golden_dir, "post", "theta.npy"))
    want_pwz = np.load(os.path.join(golden_dir, "post", "pwz.npy"))
    L = pyoracle.lib()
    th = np.zeros_like(gamma); pw = np.zeros((beta.shape[1], beta.shape[0]))
    L.orc_theta_from_gamma_raw(gamma.shape[0], gamma.shape[1], gamma, th)
    L.orc_pwz_from_beta_raw(beta.shape[0], beta.shape[1], beta, pw)
    assert np.array_equal(th, want_theta)
    assert np.max(np.abs(pw - want_pwz) / want_pwz) < 1e-15
    # and the 12-digit values the pipeline carries are these values rounded
    th12 = np.zeros_like(gamma); pw12 = np.zeros_like(pw)
    L.orc_theta_from_gamma(gamma.shape[0], gamma.shape[1], gamma, th12)
    L.orc_pwz_from_beta(beta.shape[0], beta.shape[1], beta, pw12)
    assert np.max(np.abs(th12 - want_theta) / want_theta) < 1e-11
    assert np.max(np.abs(pw12 - want_pwz) / want_pwz) < 1e-11
