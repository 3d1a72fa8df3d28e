#!/usr/bin/env python
"""Pin the ARITHMETIC of the reference's lda_post.py (theta = gamma / sum(gamma), p(w|z) =
exp(beta) / sum_w exp(beta)) by running the reference's own source lines under Python 3.

TEST INFRASTRUCTURE ONLY.  Run in the build container (where /root/reference exists):

    python oracle/pin_lda_post.py

lda_post.py is a Python-2 script (print statements :7-9, `map` used as a list :50, xrange :118,
np.str, bytes written to a text file, a tab/space mix at :63) and cannot run as a whole here; no
python2 exists in this image.  What CAN run unchanged is the arithmetic: the statements at
:45-46 (row -> float64 array, builtin sum), the list comprehension inside :50, and the block
:88-96 (exp per entry, builtin sum left to right, division, column store).  This script reads
those lines out of /root/reference/lda_post.py, checks that they still say what is expected,
executes THEM (not a restatement) on the committed golden final.gamma / final.beta of
tests/golden/c0_k20 (written by the reference engine), and stores the values at full double
precision in tests/golden/post/.  tests/test_oracle_pin.py then holds oracle/score_oracle.c to
these values, and the GPU test holds oni_lda_post(text_digits=0) to them.

What stays UNPINNED: the text layer -- Python-2 str(float) prints 12 significant digits, which
is what the Scala stage parses; Python 3 prints the shortest round-trip string instead, so the
12-digit convention is restated (oracle/score_oracle.c) and not run.
"""
import csv
import os
import re
import sys
import textwrap

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/lda_post.py"
GOLD = os.path.join(ROOT, "tests", "golden")


def main():
    src = open(REF).read().split("\n")
    line = lambda n: src[n - 1]  # noqa: E731  (1-based, as cited)
    # ---- theta (lda_post.py:35-59)
    assert line(37).strip() == "readr = csv.reader(csvfile, delimiter=' ', quotechar='\"')"
    assert line(45).strip() == "t = np.asarray(row,dtype=np.float64)"
    assert line(46).strip() == "total_t = sum(t)"
    assert line(49).strip() == "if total_t > 0:"
    m = re.match(r"\s*norm_t = map\(str, (\[.*\])\)\s*$", line(50))
    assert m, line(50)
    ratio_expr = m.group(1)  # "[ti/total_t for ti in t]"
    theta_rows = []
    with open(os.path.join(GOLD, "c0_k20", "final.gamma")) as csvfile:
        ns = {"np": np, "csv": csv, "csvfile": csvfile}
        exec(line(37).strip(), ns)
        for row in ns["readr"]:
            ns["row"] = row
            exec(line(45).strip(), ns)
            exec(line(46).strip(), ns)
            if eval(line(49).strip()[3:-1], ns):
                theta_rows.append([float(x) for x in eval(ratio_expr, ns)])
            else:
                theta_rows.append([0.0] * 20)  # :54, twenty "0.0"
    theta = np.asarray(theta_rows, dtype=np.float64)

    # ---- p(w|z) (lda_post.py:70, 88-96)
    assert line(70).strip() == "words = np.loadtxt(word_prob, np.float64)"
    assert line(88).strip() == "p_wgz = np.empty([words.shape[1], words.shape[0]], dtype=np.float64)"
    block = textwrap.dedent("\n".join(src[88:96]).replace("\t", "    "))  # :89-96
    assert "raw_w = [np.exp(wi) for wi in w]" in block and "total_w = sum(raw_w)" in block
    assert "norm_w = [rwi/total_w for rwi in raw_w]" in block and "p_wgz[:,col] = norm_w" in block
    ns = {"np": np, "word_prob": os.path.join(GOLD, "c0_k20", "final.beta")}
    exec(line(70).strip(), ns)
    exec(line(88).strip(), ns)
    exec(block, ns)
    pwz = np.ascontiguousarray(ns["p_wgz"])

    out = os.path.join(GOLD, "post")
    os.makedirs(out, exist_ok=True)
    np.save(os.path.join(out, "theta.npy"), theta)
    np.save(os.path.join(out, "pwz.npy"), pwz)
    with open(os.path.join(out, "README"), "w") as f:
        f.write("theta [D][K] and p(w|z) [V][K] at full double precision: the arithmetic lines of\n"
                "/root/reference/lda_post.py (:45-46, :50, :88-96) executed under Python %d.%d / numpy %s\n"
                "on tests/golden/c0_k20/final.gamma and final.beta.  Written by oracle/pin_lda_post.py.\n"
                % (sys.version_info[0], sys.version_info[1], np.__version__))
    print("theta %s  pwz %s  written to %s" % (theta.shape, pwz.shape, out))


if __name__ == "__main__":
    main()
