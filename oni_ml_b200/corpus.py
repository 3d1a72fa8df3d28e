"""Corpus containers and the model.dat grammar (host side of the C ABI; numpy only).

A corpus is kept as flat CSR arrays (the device layout):

    doc_ptr  int64 [D+1]   first pair of each document
    word     int32 [NNZ]   word id of each (document, word) pair
    count    int32 [NNZ]   count of each pair

``model.dat`` is the reference's text form of the same thing: one line per
document, ``N w:c w:c ...`` (oni-lda-c/lda-data.c:35-51, readme.txt:172;
written by lda_pre.py:94).  Duplicate word ids inside a document are legal and
are kept as separate pairs (lda_pre.py:66-73 appends without merging).

The synthetic corpus generator of the tests and of bench.py lives outside the
package (harness/synth.py: it needs torch, the product does not).
"""
from __future__ import annotations

import dataclasses

import numpy as np


@dataclasses.dataclass
class Corpus:
    doc_ptr: np.ndarray  # int64 [D+1]
    word: np.ndarray     # int32 [NNZ]
    count: np.ndarray    # int32 [NNZ]
    num_terms: int

    @property
    def num_docs(self) -> int:
        return int(self.doc_ptr.shape[0] - 1)

    @property
    def nnz(self) -> int:
        return int(self.word.shape[0])

    @property
    def tokens(self) -> int:
        return int(self.count.astype(np.int64).sum())

    def slice_docs(self, lo: int, hi: int) -> "Corpus":
        """Documents [lo, hi) as a corpus of their own (same vocabulary)."""
        a, b = int(self.doc_ptr[lo]), int(self.doc_ptr[hi])
        return Corpus((self.doc_ptr[lo:hi + 1] - a).astype(np.int64),
                      self.word[a:b].copy(), self.count[a:b].copy(), self.num_terms)


def read_model_dat(path: str) -> Corpus:
    """Python mirror of read_data (oni-lda-c/lda-data.c:22-59), for tests."""
    ptr = [0]
    words: list[int] = []
    counts: list[int] = []
    with open(path, "r") as f:
        toks = f.read().split()
    i = 0
    while i < len(toks):
        n = int(toks[i]); i += 1
        for _ in range(n):
            w, c = toks[i].split(":"); i += 1
            words.append(int(w)); counts.append(int(c))
        ptr.append(len(words))
    word = np.asarray(words, dtype=np.int32)
    nt = int(word.max()) + 1 if word.size else 0
    return Corpus(np.asarray(ptr, dtype=np.int64), word,
                  np.asarray(counts, dtype=np.int32), nt)


def write_model_dat(path: str, c: Corpus) -> None:
    """Write the reference's model.dat text (lda_pre.py:94: count, then ' w:c'...)."""
    ptr, word, count = c.doc_ptr, c.word, c.count
    with open(path, "w") as f:
        lines = []
        for d in range(c.num_docs):
            a, b = int(ptr[d]), int(ptr[d + 1])
            lines.append(str(b - a) + "".join(" %d:%d" % (word[j], count[j])
                                              for j in range(a, b)) + "\n")
            if len(lines) >= 4096:
                f.write("".join(lines))
                lines = []
        f.write("".join(lines))
