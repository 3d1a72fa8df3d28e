#!/usr/bin/env python
"""Generate the committed golden fixtures under tests/golden/ by RUNNING THE
REFERENCE ITSELF (oracle/_ref/lda_ref = the unmodified oni-lda-c sources built
by oracle/ref_build/Makefile) on small synthetic corpora.

TEST INFRASTRUCTURE ONLY.  Run from the repo root, in the build container
(where /root/reference exists):

    make -C oracle && python oracle/make_golden.py

The reference has no tests or golden vectors of its own,
so these outputs are the parity pins: 000.beta pins the RNG, the random
initialisation and lda_mle; likelihood.dat line i pins the E-step at EM
iteration i (line 1 needs only the initialisation and one E-step) and, from
line 2 on, the reduction, M-step and alpha Newton; final.* pin the whole run.
"""
import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from harness.synth import synth_corpus
from oni_ml_b200.corpus import Corpus, write_model_dat  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref", "lda_ref")
GOLD = os.path.join(ROOT, "tests", "golden")

SETTINGS_SHIPPED = ("var max iter 20\nvar convergence 1e-6\nem max iter 20\n"
                    "em convergence 1e-6\nalpha estimate\n")  # oni-lda-c/settings.txt
SETTINGS_FIXED = ("var max iter 5\nvar convergence 1e-4\nem max iter 6\n"
                  "em convergence 1e-5\nalpha fixed\n")
SETTINGS_UNBOUNDED = ("var max iter -1\nvar convergence 1e-6\nem max iter 4\n"
                      "em convergence 1e-4\nalpha estimate\n")  # cf. inf-settings.txt


def run_ref(model, settings_text, alpha, k, seed, nproc, keep, outdir, start="random"):
    with tempfile.TemporaryDirectory() as tmp:
        st = os.path.join(tmp, "settings.txt")
        with open(st, "w") as f:
            f.write(settings_text)
        out = os.path.join(tmp, "out")
        env = dict(os.environ, ONI_LDA_SEED=str(seed), ONI_MPI_NP=str(nproc))
        subprocess.run([REF, "est", alpha, str(k), st, str(nproc), model, start, out],
                       check=True, env=env, stdout=subprocess.DEVNULL)
        os.makedirs(outdir, exist_ok=True)
        with open(os.path.join(outdir, "settings.txt"), "w") as f:
            f.write(settings_text)
        with open(os.path.join(outdir, "args.txt"), "w") as f:
            f.write("alpha %s\nk %d\nseed %d\nnproc %d\n" % (alpha, k, seed, nproc))
        for name in keep:
            shutil.copy(os.path.join(out, name), os.path.join(outdir, name))


def make_c1():
    """BASELINE.json configs[0]: 10k IP documents, 5k-word vocabulary, K=20, alpha=2.5, one
    rank.  The corpus is regenerated from its seed by the tests (too large to commit as
    text); kept here are likelihood.dat, final.other, 000.other and every 50th row of
    final.gamma / every 25th column of final.beta of the reference's run."""
    c = synth_corpus(10000, 5000, k_true=20, mu=2.5, seed=20160123)
    out = os.path.join(GOLD, "c1_k20")
    with tempfile.TemporaryDirectory() as tmp:
        model = os.path.join(tmp, "model.dat")
        write_model_dat(model, c)
        run_ref(model, SETTINGS_SHIPPED, "2.5", 20, 4357, 1,
                ["000.other", "likelihood.dat", "final.other", "final.gamma", "final.beta"], out)
    gam = np.loadtxt(os.path.join(out, "final.gamma"))
    beta = np.loadtxt(os.path.join(out, "final.beta"))
    np.savetxt(os.path.join(out, "final.gamma.rows50"), gam[::50], fmt="%.10f")
    np.savetxt(os.path.join(out, "final.beta.cols25"), beta[:, ::25], fmt="%.10f")
    # the whole of both files as the doubles their 10-decimal text parses to (for the chain-level
    # ranked-list test: reference EM -> lda_post.py shaping -> scoring, beside the GPU chain)
    np.savez_compressed(os.path.join(out, "final_full.npz"), gamma=gam, beta=beta)
    os.remove(os.path.join(out, "final.gamma"))
    os.remove(os.path.join(out, "final.beta"))
    with open(os.path.join(out, "corpus.txt"), "w") as f:
        f.write("synth_corpus(10000, 5000, k_true=20, mu=2.5, seed=20160123): %d docs %d pairs %d tokens\n"
                % (c.num_docs, c.nnz, c.tokens))
        f.write("crc word %d count %d ptr %d\n" % (int(c.word.astype(np.int64).sum()),
                                                   int(c.count.astype(np.int64).sum()), int(c.doc_ptr.sum())))


def make_c2(nproc=None):
    """BASELINE.json configs[1], the size the headline is quoted on: 1M documents, 50k words,
    K=20, shipped settings, `est 2.5 20 ... random` (ml_ops.sh:80) on the reference binary with
    `nproc` forked ranks (lda-estimate.c:201 deals documents d = myid, myid+nproc, ...).  Takes
    the better part of an hour on 8 cores, so it is not part of main(): run
    `python oracle/make_golden.py c2`.  Kept: likelihood.dat, final.other, 000.other, every
    5000th row of final.gamma, every 250th column of final.beta, and a CRC of the corpus the
    tests regenerate from its seed (CPU generator: the same stream on every machine)."""
    nproc = nproc or max(1, (os.cpu_count() or 2) - 1)
    c = synth_corpus(1_000_000, 50_000, k_true=20, mu=3.5, seed=20160123)
    out = os.path.join(GOLD, "c2_flow_day")
    with tempfile.TemporaryDirectory() as tmp:
        model = os.path.join(tmp, "model.dat")
        write_model_dat(model, c)
        run_ref(model, SETTINGS_SHIPPED, "2.5", 20, 4357, nproc,
                ["000.other", "likelihood.dat", "final.other", "final.gamma", "final.beta"], out)
    gam = np.loadtxt(os.path.join(out, "final.gamma"))
    beta = np.loadtxt(os.path.join(out, "final.beta"))
    np.savetxt(os.path.join(out, "final.gamma.rows5000"), gam[::5000], fmt="%.10f")
    np.savetxt(os.path.join(out, "final.beta.cols250"), beta[:, ::250], fmt="%.10f")
    os.remove(os.path.join(out, "final.gamma"))
    os.remove(os.path.join(out, "final.beta"))
    with open(os.path.join(out, "corpus.txt"), "w") as f:
        f.write("synth_corpus(1000000, 50000, k_true=20, mu=3.5, seed=20160123): %d docs %d pairs %d tokens\n"
                % (c.num_docs, c.nnz, c.tokens))
        f.write("crc word %d count %d ptr %d\n" % (int(c.word.astype(np.int64).sum()),
                                                   int(c.count.astype(np.int64).sum()), int(c.doc_ptr.sum())))


def make_pre():
    """lda_pre.py fixture: a doc_wc.dat shaped like flow_pre_lda.scala's output (ip,word,count;
    words repeat inside a document, one field is quoted) and what the REFERENCE script
    (/root/reference/lda_pre.py, runs unmodified under Python 3) writes for it."""
    import random
    rnd = random.Random(5)
    out = os.path.join(GOLD, "pre")
    os.makedirs(out, exist_ok=True)
    words = ["%d.0_%d.0_%d.0_%d.0" % (rnd.randint(0, 65535), rnd.randint(0, 9), rnd.randint(0, 9), rnd.randint(0, 9))
             for _ in range(120)] + ["-1_80.0_3.0_2.0_1.0", '"quoted,word"']
    ips = ["10.%d.%d.%d" % (rnd.randint(0, 3), rnd.randint(0, 255), rnd.randint(0, 255)) for _ in range(80)]
    rows = ["%s,%s,%d" % (rnd.choice(ips), rnd.choice(words), rnd.randint(1, 900)) for _ in range(1500)]
    rows.append("10.9.9.9,solo_word,7")
    with open(os.path.join(out, "doc_wc.dat"), "w") as f:
        f.write("\n".join(rows) + "\n")
    subprocess.run([sys.executable, "/root/reference/lda_pre.py", out + "/"], check=True, stdout=subprocess.DEVNULL)


def main():
    os.makedirs(GOLD, exist_ok=True)
    c = synth_corpus(400, 300, k_true=8, mu=2.0, seed=20160122)
    # edge cases the grammar allows: an empty document, a one-word document and
    # a document made of one repeated word id (kept as separate pairs)
    ptr = c.doc_ptr.tolist()
    word = c.word.tolist()
    count = c.count.tolist()
    for w_list, c_list in (([], []), ([7], [3]), ([11, 11, 11], [1, 2, 5])):
        word += w_list
        count += c_list
        ptr.append(len(word))
    c = Corpus(np.asarray(ptr, np.int64), np.asarray(word, np.int32),
               np.asarray(count, np.int32), 300)
    model = os.path.join(GOLD, "c0_model.dat")
    write_model_dat(model, c)

    everything = ["000.beta", "000.other", "likelihood.dat", "final.beta", "final.gamma",
                  "final.other"]
    run_ref(model, SETTINGS_SHIPPED, "2.5", 20, 4357, 1, everything,
            os.path.join(GOLD, "c0_k20"))
    run_ref(model, SETTINGS_SHIPPED, "2.5", 20, 4357, 3,
            ["likelihood.dat", "final.other"], os.path.join(GOLD, "c0_k20_np3"))
    run_ref(model, SETTINGS_FIXED, "0.1", 5, 99, 1,
            ["000.other", "likelihood.dat", "final.gamma", "final.other"],
            os.path.join(GOLD, "c0_k5_fixed"))
    run_ref(model, SETTINGS_UNBOUNDED, "1.0", 37, 12345, 1,
            ["likelihood.dat", "final.other"], os.path.join(GOLD, "c0_k37_unbounded"))
    # the other two starts of lda-estimate.c:147-167 and `lda inf` (lda-estimate.c:358-388)
    short = ("var max iter 20\nvar convergence 1e-6\nem max iter 3\nem convergence 1e-6\nalpha estimate\n")
    run_ref(model, short, "2.5", 20, 777, 1, ["000.beta", "likelihood.dat", "final.other"],
            os.path.join(GOLD, "c0_k20_seeded"), start="seeded")
    run_ref(model, short, "2.5", 20, 1, 1, ["000.other", "likelihood.dat", "final.other"],
            os.path.join(GOLD, "c0_k20_resume"), start=os.path.join(GOLD, "c0_k20", "final"))
    inf_dir = os.path.join(GOLD, "c0_k20_inf")
    os.makedirs(inf_dir, exist_ok=True)
    with open(os.path.join(inf_dir, "settings.txt"), "w") as f:
        f.write("var max iter -1\nvar convergence 1e-6\nem max iter 100\nem convergence 1e-4\nalpha estimate\n")
    with tempfile.TemporaryDirectory() as tmp:
        subprocess.run([REF, "inf", os.path.join(inf_dir, "settings.txt"), os.path.join(GOLD, "c0_k20", "final"),
                        model, os.path.join(tmp, "c0")], check=True, stdout=subprocess.DEVNULL,
                       env=dict(os.environ, ONI_MPI_NP="1"))
        shutil.copy(os.path.join(tmp, "c0-lda-lhood.dat"), inf_dir)
        shutil.copy(os.path.join(tmp, "c0-gamma.dat"), inf_dir)
    make_c1()
    make_pre()
    print("golden fixtures written to", GOLD)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "c2":
        make_c2(int(sys.argv[2]) if len(sys.argv) > 2 else None)
    elif len(sys.argv) > 1 and sys.argv[1] == "c1":
        make_c1()
    else:
        main()
