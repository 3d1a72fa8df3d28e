"""Documents sharded over several GPUs of one box, one NCCL all-reduce of the packed
sufficient statistics per EM iteration (replaces the 2K+3 MPI_Allreduce calls of
oni-lda-c/lda-estimate.c:211-252).  Needs >= 2 devices; skipped otherwise."""
import os

import numpy as np
import pytest

from oni_ml_b200 import files, lda

pytestmark = pytest.mark.gpu


def _ndev():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("ngpus", [2, 4, 8])
def test_sharded_em_matches_single_gpu_and_reference(golden_dir, c0_corpus, ngpus):
    if _ndev() < ngpus:
        pytest.skip("needs %d GPUs" % ngpus)
    d = os.path.join(golden_dir, "c0_k20")
    st = lda.read_settings(os.path.join(d, "settings.txt"))
    one = lda.estimate(c0_corpus, 20, 2.5, st, 4357, ngpus=1)
    many = lda.estimate(c0_corpus, 20, 2.5, st, 4357, ngpus=ngpus)
    # only the order of the cross-shard sums differs
    assert many["iters"] == one["iters"] == 21
    assert np.allclose(many["lik"], one["lik"], rtol=1e-10)
    assert np.allclose(many["gamma"], one["gamma"], rtol=1e-7)
    assert abs(many["alpha"] - one["alpha"]) < 1e-9 * one["alpha"]
    want = files.read_likelihood(os.path.join(d, "likelihood.dat"))
    assert np.max(np.abs(many["lik"] - want[:, 0]) / np.abs(want[:, 0])) < 1e-6
    gam = files.read_matrix(os.path.join(d, "final.gamma"))
    assert np.max(np.abs(many["gamma"] - gam) / (np.abs(gam) + 1e-4)) < 1e-6


def test_cli_with_two_workers(golden_dir, tmp_path):
    if _ndev() < 2:
        pytest.skip("needs 2 GPUs")
    d = os.path.join(golden_dir, "c0_k20")
    out = tmp_path / "day"
    p = lda.run_cli(["est", "2.5", "20", os.path.join(d, "settings.txt"), "2",
                     os.path.join(golden_dir, "c0_model.dat"), "random", str(out)], seed=4357)
    assert "on 2 GPU(s)" in p.stdout
    lik = files.read_likelihood(str(out / "likelihood.dat"))
    want = files.read_likelihood(os.path.join(d, "likelihood.dat"))
    assert np.max(np.abs(lik[:, 0] - want[:, 0]) / np.abs(want[:, 0])) < 1e-6
    assert open(out / "000.beta").read() == open(os.path.join(d, "000.beta")).read()


@pytest.mark.parametrize("ngpus", [1, 2, 4, 8])
def test_scoring_over_several_gpus_is_the_single_gpu_list(ngpus):
    """oni_score_flow_multi / oni_score_dns_multi: events cut into contiguous shares, tables
    replicated, ranked shares merged on (key, event index) -- the same list, bit for bit, with
    ties (events repeated on purpose) landing in index order across the share boundaries."""
    if _ndev() < ngpus:
        pytest.skip("needs %d GPUs" % ngpus)
    from oni_ml_b200 import scoring
    rng = np.random.default_rng(3)
    D, V, K, n = 3000, 700, 20, 200_001
    theta = rng.dirichlet(np.full(K, 0.3), D)
    pwz = rng.dirichlet(np.full(V, 0.05), K).T.copy()
    ev = [rng.integers(-1, lim, n).astype(np.int32) for lim in (D, D, V, V)]
    for a in ev:  # identical events in different shares: equal keys
        a[n // 2:n // 2 + 5000] = a[:5000]
    one = scoring.score_flow(theta, pwz, *ev, 1.0)
    tol = float(np.quantile(one["key"], 0.02))
    for t, cap in ((1.0, None), (tol, None), (tol, 777)):
        a = scoring.score_flow(theta, pwz, *ev, t, out_cap=cap)
        b = scoring.score_flow(theta, pwz, *ev, t, out_cap=cap, ngpus=ngpus)
        assert a["n"] == b["n"] and np.array_equal(a["idx"], b["idx"])
        for name in ("key", "src", "dst"):
            assert np.array_equal(a[name].view(np.int64), b[name].view(np.int64)), name
    a = scoring.score_dns(theta, pwz, ev[0], ev[2], tol)
    b = scoring.score_dns(theta, pwz, ev[0], ev[2], tol, ngpus=ngpus)
    assert a["n"] == b["n"] and np.array_equal(a["idx"], b["idx"])
    assert np.array_equal(a["key"].view(np.int64), b["key"].view(np.int64))
    with pytest.raises(Exception):
        scoring.score_flow(theta, pwz, *ev, tol, ngpus=_ndev() + 1)
