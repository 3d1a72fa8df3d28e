"""The synthetic corpus generator of the tests, bench.py and oracle/make_golden.py.

Follows SURVEY.md section 8(d): LDA generative model with Zipf-shaped topics, log-normal
document lengths, geometric counts, a few documents with one repeated word id, and every word
id present at least once so that ``num_terms == V`` (lda-data.c:49 takes max id + 1).  Uses
torch's generators (CPU: the same stream on every machine; CUDA: a flow day in HBM without a host
round trip) -- harness code, kept out of the product package.
"""
from __future__ import annotations

import numpy as np
import torch

from oni_ml_b200.corpus import Corpus


def synth_corpus(num_docs: int, num_terms: int, k_true: int = 20, mu: float = 2.5,
                 sigma: float = 1.0, geom_p: float = 0.25, dup_frac: float = 0.02,
                 seed: int = 20160122, device: str | torch.device = "cpu",
                 as_torch: bool = False, topic_seed: int | None = None):
    """Synthetic corpus of SURVEY.md section 8(d).

    Returns a :class:`Corpus` (numpy, host) or, with ``as_torch=True``, the three
    tensors ``(doc_ptr int64, word int32, count int32)`` on ``device`` so a
    flow-day sized corpus can be generated in HBM without a host round trip.

    ``topic_seed`` draws the topics from their own stream, so that the shards of a
    multi-GPU run (different ``seed`` per rank) are documents of ONE corpus: same
    topics, different documents.  ``None`` keeps the single-stream behaviour the
    committed fixtures were generated with.
    """
    dev = torch.device(device)
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    D, V, K = int(num_docs), int(num_terms), int(k_true)
    f64 = torch.float64

    def gamma_sample(shape, conc):
        # Gamma(conc, 1) via torch._standard_gamma (generator-aware)
        return torch._standard_gamma(torch.full(shape, conc, dtype=f64, device=dev), generator=g)

    # topics: Dirichlet(0.05) shaped by a Zipf(1.05) envelope over a permuted vocabulary
    zipf = 1.0 / torch.arange(1, V + 1, dtype=f64, device=dev) ** 1.05
    g_docs = g
    if topic_seed is not None:
        g = torch.Generator(device=dev)
        g.manual_seed(topic_seed)
    perm = torch.randperm(V, generator=g, device=dev)
    beta = gamma_sample((K, V), 0.05) * zipf[perm][None, :] + 1e-300
    g = g_docs
    beta /= beta.sum(1, keepdim=True)
    cdf = torch.cumsum(beta, 1)
    cdf = cdf / cdf[:, -1:]
    cdf_flat = (cdf + torch.arange(K, dtype=f64, device=dev)[:, None]).reshape(-1)

    theta = gamma_sample((D, K), 0.5) + 1e-300
    theta /= theta.sum(1, keepdim=True)
    theta_cdf = torch.cumsum(theta, 1)

    n_d = torch.exp(mu + sigma * torch.randn(D, generator=g, device=dev, dtype=f64))
    n_d = n_d.round().clamp_(1, V).to(torch.int64)
    ptr = torch.zeros(D + 1, dtype=torch.int64, device=dev)
    torch.cumsum(n_d, 0, out=ptr[1:])
    nnz = int(ptr[-1])

    word = torch.empty(nnz, dtype=torch.int32, device=dev)
    doc_of = torch.repeat_interleave(torch.arange(D, device=dev), n_d)
    chunk = 1 << 22
    for a in range(0, nnz, chunk):
        b = min(nnz, a + chunk)
        u = torch.rand(b - a, generator=g, device=dev, dtype=f64)
        z = (u[:, None] > theta_cdf[doc_of[a:b]]).sum(1).clamp_(max=K - 1)
        u2 = torch.rand(b - a, generator=g, device=dev, dtype=f64)
        pos = torch.searchsorted(cdf_flat, z.to(f64) + u2)
        word[a:b] = (pos - z * V).clamp_(0, V - 1).to(torch.int32)
    # geometric counts >= 1
    u3 = torch.rand(nnz, generator=g, device=dev, dtype=f64).clamp_(min=1e-300)
    count = (torch.floor(torch.log(u3) / np.log1p(-geom_p)) + 1).clamp_(1, 1 << 20).to(torch.int32)

    # a few documents repeat one word id (flow quirk, flow_pre_lda.scala:366-373)
    if dup_frac > 0:
        cand = torch.nonzero(n_d >= 2).flatten()
        ndup = int(dup_frac * D)
        if ndup and cand.numel():
            pick = cand[torch.randperm(cand.numel(), generator=g, device=dev)[:ndup]]
            word[ptr[pick + 1] - 1] = word[ptr[pick]]

    # every word id must appear so that num_terms == V
    seen = torch.zeros(V, dtype=torch.bool, device=dev)
    seen[word.long()] = True
    missing = torch.nonzero(~seen).flatten()
    if missing.numel():
        m = missing.numel()
        extra_doc = torch.arange(m, device=dev) % D
        all_doc = torch.cat([doc_of, extra_doc])
        all_word = torch.cat([word, missing.to(torch.int32)])
        all_count = torch.cat([count, torch.ones(m, dtype=torch.int32, device=dev)])
        order = torch.sort(all_doc, stable=True).indices
        word, count = all_word[order], all_count[order]
        n_d = n_d + torch.bincount(extra_doc, minlength=D)
        torch.cumsum(n_d, 0, out=ptr[1:])

    if as_torch:
        return ptr, word.contiguous(), count.contiguous()
    return Corpus(ptr.cpu().numpy(), word.cpu().numpy(), count.cpu().numpy(), V)
