"""Python view of the engine's output files (the file contract).

The product writes these files from C (csrc/lda_io.c); this module formats and
parses the same grammar so that tests can compare arrays with files and files
with the reference's bytes.

    <root>.beta     K lines, every value printed " %5.10f" (leading blank)
                    oni-lda-c/lda-model.c:199-210
    <root>.other    "num_topics %d\\nnum_terms %d\\nalpha %5.10f\\n"   :212-216
    final.gamma     D lines, "%5.10f" joined by one blank   lda-estimate.c:88-104
    likelihood.dat  "%10.10f\\t%5.5e\\n" per EM iteration   lda-estimate.c:266
"""
from __future__ import annotations

import numpy as np


def format_beta(logbeta: np.ndarray) -> str:
    """logbeta [K][V] (topic-major) -> text of <root>.beta."""
    return "".join("".join(" %5.10f" % v for v in row) + "\n" for row in logbeta)


def format_other(k: int, v: int, alpha: float) -> str:
    return "num_topics %d\nnum_terms %d\nalpha %5.10f\n" % (k, v, alpha)


def format_gamma(gamma: np.ndarray) -> str:
    return "".join(" ".join("%5.10f" % v for v in row) + "\n" for row in gamma)


def format_likelihood(lik, conv) -> str:
    return "".join("%10.10f\t%5.5e\n" % (l, c) for l, c in zip(lik, conv))


def read_matrix(path: str) -> np.ndarray:
    """final.beta / final.gamma as float64 [rows][cols]."""
    return np.atleast_2d(np.loadtxt(path, dtype=np.float64))


def read_other(path: str) -> dict:
    out = {}
    with open(path) as f:
        for line in f:
            key, val = line.split()
            out[key] = float(val) if key == "alpha" else int(val)
    return out


def read_likelihood(path: str) -> np.ndarray:
    """likelihood.dat -> float64 [iters][2] (bound, relative change; first change is inf)."""
    rows = []
    with open(path) as f:
        for line in f:
            a, b = line.split()
            rows.append((float(a), float(b)))
    return np.asarray(rows, dtype=np.float64)
