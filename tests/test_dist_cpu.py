"""The N > 1 host path on CPU: world_size-2 `gloo` ranks drive oni_ml_b200.dist
(document sharding + the EM loop with ONE all-reduce of the packed statistics
per iteration).  The per-rank arithmetic is supplied by the oracle here (tests
may use it as a stand-in; there is no GPU in this container) -- what is under
test is the sharding, the packing and the reduce-then-replicated-M-step
structure, against the single-rank oracle run."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oni_ml_b200 import dist as odist
from oni_ml_b200.corpus import read_model_dat
from oracle import pyoracle

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
K, ALPHA, SEED = 20, 2.5, 4357


class OracleRank:
    """Same step surface as lda.LdaEngine, arithmetic from the oracle."""

    def __init__(self, shard, num_terms, st, logbeta0, alpha):
        self.shard, self.V, self.st = shard, num_terms, st
        self.logbeta, self.alpha_v = logbeta0.copy(), alpha
        self.var_max_iter = st.var_max_iter
        self.em_iter, self.lik_old, self.converged = 0, 0.0, 1.0
        self.packed = None
        self.gamma_v = None

    def e_step(self):
        r = pyoracle.e_step(self.shard, K, self.alpha_v, self.logbeta, self.var_max_iter,
                            self.st.var_converged)
        self.gamma_v = r["gamma"]
        # [V*K class_word word-major | K class_total | alpha_ss, likelihood, num_docs, visits]
        self.packed = np.concatenate([r["class_word"].T.ravel(), r["class_total"],
                                      [r["alpha_ss"], r["lik"], r["num_docs"], r["word_visits"]]])

    def m_step(self):
        L = pyoracle.lib()
        V = self.V
        cw = np.ascontiguousarray(self.packed[:V * K].reshape(V, K).T)
        ct = np.ascontiguousarray(self.packed[V * K:V * K + K])
        ass, lik, nd = self.packed[V * K + K:V * K + K + 3]
        L.orc_mle_beta(K, V, cw, ct, self.logbeta)
        if self.st.estimate_alpha == 1:
            self.alpha_v = L.orc_opt_alpha(float(ass), int(nd), K)
        self.em_iter += 1
        with np.errstate(divide="ignore", invalid="ignore"):
            self.converged = float((np.float64(self.lik_old) - lik) / np.float64(self.lik_old))
        if self.converged < 0 and self.var_max_iter > 0:
            self.var_max_iter *= 2
        self.lik_old = float(lik)
        return float(lik), self.converged

    def should_continue(self):
        return ((self.converged < 0 or self.converged > self.st.em_converged or self.em_iter <= 2)
                and self.em_iter <= self.st.em_max_iter)


def _worker(rank, world, port, outdir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    c = read_model_dat(os.path.join(GOLDEN, "c0_model.dat"))
    st = pyoracle.Settings(20, 1e-6, 5, 1e-6, 1)
    cut = odist.shard_bounds(c.doc_ptr, world)
    shard = c.slice_docs(cut[rank], cut[rank + 1])
    L = pyoracle.lib()
    L.orc_mt_seed(SEED)
    cw = np.zeros((K, c.num_terms)); ct = np.zeros(K); lb0 = np.zeros((K, c.num_terms))
    L.orc_random_init_ss(K, c.num_terms, cw, ct)
    L.orc_mle_beta(K, c.num_terms, cw, ct, lb0)
    eng = OracleRank(shard, c.num_terms, st, lb0, ALPHA)

    def allreduce():
        t = torch.from_numpy(eng.packed)
        dist.all_reduce(t)  # one collective per EM iteration

    lik, conv = odist.run_em(eng, allreduce)
    np.savez(os.path.join(outdir, "rank%d.npz" % rank), lik=lik, conv=conv, gamma=eng.gamma_v,
             alpha=eng.alpha_v, lo=cut[rank], hi=cut[rank + 1], logbeta=eng.logbeta)
    dist.destroy_process_group()


def test_two_rank_em_matches_single_rank(tmp_path):
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    c = read_model_dat(os.path.join(GOLDEN, "c0_model.dat"))
    st = pyoracle.Settings(20, 1e-6, 5, 1e-6, 1)
    want = pyoracle.run_em(c, K, ALPHA, st, SEED)
    r0, r1 = np.load(tmp_path / "rank0.npz"), np.load(tmp_path / "rank1.npz")
    assert r0["lo"] == 0 and r0["hi"] == r1["lo"] and r1["hi"] == c.num_docs
    assert len(r0["lik"]) == want["iters"] == 6
    # both ranks hold the same replicated model and trajectory
    assert np.array_equal(r0["lik"], r1["lik"]) and np.array_equal(r0["logbeta"], r1["logbeta"])
    assert np.allclose(r0["lik"], want["lik"], rtol=1e-10)
    assert abs(float(r0["alpha"]) - want["alpha"]) < 1e-9 * want["alpha"]
    gam = np.vstack([r0["gamma"], r1["gamma"]])
    assert np.allclose(gam, want["gamma"], rtol=1e-7)


@pytest.mark.parametrize("n", [1, 2, 3, 8])
def test_shard_bounds_cover_and_balance(n):
    c = read_model_dat(os.path.join(GOLDEN, "c0_model.dat"))
    cut = odist.shard_bounds(c.doc_ptr, n)
    assert cut[0] == 0 and cut[-1] == c.num_docs and all(a <= b for a, b in zip(cut, cut[1:]))
    pairs = [int(c.doc_ptr[b] - c.doc_ptr[a]) for a, b in zip(cut, cut[1:])]
    assert sum(pairs) == c.nnz
    longest = int(np.diff(c.doc_ptr).max())
    assert max(pairs) - min(pairs) <= 2 * longest + 1


def test_shard_bounds_more_ranks_than_documents():
    ptr = np.array([0, 3, 3, 10], dtype=np.int64)
    cut = odist.shard_bounds(ptr, 8)
    assert cut[0] == 0 and cut[-1] == 3 and all(a <= b for a, b in zip(cut, cut[1:]))
