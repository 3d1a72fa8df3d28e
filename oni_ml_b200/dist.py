"""Document sharding and the multi-rank EM driver (host logic, no arithmetic).

The reference shards documents round-robin over MPI ranks and reduces the
sufficient statistics with 2K+3 MPI_Allreduce calls per EM iteration
(oni-lda-c/lda-estimate.c:201,211-252).  Here every rank (= one process = one
GPU) owns a contiguous, pair-balanced block of documents and the packed
statistics are summed by ONE all-reduce per iteration: the library's
ncclAllReduce on the engine's own stream.  How the 128-byte NCCL id reaches the
ranks is the launcher's business (harness/torchrun.py does it with
torch.distributed for the tests and bench.py).
"""
from __future__ import annotations

from typing import Callable

import numpy as np


def shard_bounds(doc_ptr: np.ndarray, nranks: int) -> list[int]:
    """Cut points cut[r]..cut[r+1] of contiguous shards balanced by pair count
    (same rule as shard_docs in csrc/engine.cu)."""
    d = int(doc_ptr.shape[0]) - 1
    nnz = int(doc_ptr[-1])
    cut = [0] * (nranks + 1)
    cut[nranks] = d
    for r in range(1, nranks):
        target = nnz * r // nranks
        pos = int(np.searchsorted(doc_ptr, target, side="left"))
        cut[r] = max(cut[r - 1], min(pos, d))
    return cut


def em_iteration(engine, allreduce: Callable[[], None] | None = None):
    """One EM iteration of one rank: E-step over the shard, one all-reduce of the
    packed statistics, M-step (replicated).  `allreduce` defaults to the
    engine's own (NCCL) collective."""
    engine.e_step()
    (allreduce or engine.allreduce)()
    return engine.m_step()


def run_em(engine, allreduce: Callable[[], None] | None = None, on_iter=None):
    """The EM loop of lda-estimate.c:186-270 for one rank; returns (lik, conv) lists."""
    lik, conv = [], []
    while engine.should_continue():
        l, c = em_iteration(engine, allreduce)
        lik.append(l)
        conv.append(c)
        if on_iter:
            on_iter(len(lik), l, c)
    return lik, conv
