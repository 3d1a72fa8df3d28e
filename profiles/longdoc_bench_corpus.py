import os, sys
import numpy as np
sys.path.insert(0, os.getcwd())
os.environ["ONI_LDA_TIMING"] = "1"
import torch
from oni_ml_b200 import _lib, lda
from harness.synth import synth_corpus
V, K = 50000, 20
ptr, word, count = synth_corpus(1_000_000, V, k_true=20, mu=3.5, geom_p=0.25, seed=20160123, device="cuda", as_torch=True, topic_seed=20160123)
lens = (ptr[1:] - ptr[:-1]).cpu().numpy()
top = np.argsort(lens)[-8:][::-1]
print("longest docs", [(int(d), int(lens[d])) for d in top], flush=True)
st = _lib.Settings(20, 1e-6, 10000, 1e-6, 1)
for excl in (None, "7000", "0"):
    if excl is None: os.environ.pop("ONI_LDA_EXCL_DOC", None)
    else: os.environ["ONI_LDA_EXCL_DOC"] = excl
    print("== EXCL", excl, file=sys.stderr, flush=True)
    with lda.LdaEngine(0, K, V, st) as eng:
        eng.set_corpus_device(ptr, word, count); eng.init_random(20160123, 2.5)
        for it in range(3):
            eng.em_iteration()
        s = eng.stats(); it_ = eng.doc_iters()
        print("EXCL", excl, "E-step %.2f ms" % s["ms_estep_kernel"], "sweeps of the longest:", [int(it_[d]) for d in top], flush=True)
