"""One 8000-pair document next to a flow-day-like shard: how long does its serial chain take when
the rest of the GPU is busy?  Prints the library's ONI_LDA_TIMING line (when each launch ended)."""
import os, sys
import numpy as np
sys.path.insert(0, os.getcwd())
os.environ["ONI_LDA_TIMING"] = "1"
import torch
from oni_ml_b200 import _lib, lda
from harness.synth import synth_corpus
from oni_ml_b200.corpus import Corpus

V, K, N = 50000, 20, 8000
bg = synth_corpus(int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000, V, mu=3.5, seed=20160123)  # max doc 3657
rng = np.random.default_rng(0)
lw = rng.integers(0, V, size=N).astype(np.int32); lc = rng.geometric(0.25, size=N).astype(np.int32)
both = Corpus(np.concatenate([bg.doc_ptr, [bg.doc_ptr[-1] + N]]).astype(np.int64), np.concatenate([bg.word, lw]),
              np.concatenate([bg.count, lc]), V)
lone = Corpus(np.array([0, N], np.int64), lw, lc, V)
st = _lib.Settings(20, 1e-6, 20, 1e-6, 1)
for name, corpus, excl in (("background only", bg, None), ("long document only", lone, None),
                           ("both, shared SMs", both, "0"), ("both, long one on its own SM", both, None)):
    if excl is None:
        os.environ.pop("ONI_LDA_EXCL_DOC", None)
    else:
        os.environ["ONI_LDA_EXCL_DOC"] = excl
    print("==", name, file=sys.stderr, flush=True)
    with lda.LdaEngine(0, K, V, st) as eng:
        eng.set_corpus(corpus); eng.init_random(1, 2.5)
        for _ in range(3):
            eng.em_iteration()
            s = eng.stats()
        print("%-32s E-step %.2f ms, %d visits" % (name, s["ms_estep_kernel"], s["word_visits"]), flush=True)
