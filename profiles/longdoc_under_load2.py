"""One long document next to a shard of the size one of 8 GPUs holds of a flow day: the pace of its
serial chain alone and beside the bulk (time from the start of the E-step to the end of the
exclusive launch / visits of the document)."""
import os, sys
import numpy as np
sys.path.insert(0, os.getcwd())
from oni_ml_b200 import _lib, lda
from harness.synth import synth_corpus
from oni_ml_b200.corpus import Corpus

V, K, N = 50000, 20, int(sys.argv[2]) if len(sys.argv) > 2 else 8000
bg = synth_corpus(int(sys.argv[1]) if len(sys.argv) > 1 else 125_000, V, mu=3.5, seed=20160123)
rng = np.random.default_rng(0)
lw = rng.integers(0, V, size=N).astype(np.int32); lc = rng.geometric(0.25, size=N).astype(np.int32)
both = Corpus(np.concatenate([bg.doc_ptr, [bg.doc_ptr[-1] + N]]).astype(np.int64), np.concatenate([bg.word, lw]),
              np.concatenate([bg.count, lc]), V)
lone = Corpus(np.array([0, N], np.int64), lw, lc, V)
st = _lib.Settings(20, 1e-6, 20, 1e-6, 1)
for name, corpus in (("long document alone", lone), ("beside %d documents" % bg.num_docs, both)):
    with lda.LdaEngine(0, K, V, st) as eng:
        eng.set_corpus(corpus); eng.init_random(1, 2.5)
        for it in range(4):
            eng.em_iteration()
            s = eng.stats()
            sweeps = int(eng.doc_iters()[-1])
            print("%-28s EM %d: exclusive launch (%d docs) ended at %.2f ms, main at %.2f; the long document: %d sweeps "
                  "-> %.3f us per visit" % (name, it + 1, s["docs_exclusive"], s["ms_end_exclusive"], s["ms_end_main"],
                                            sweeps, s["ms_end_exclusive"] * 1e3 / (N * sweeps)), flush=True)
