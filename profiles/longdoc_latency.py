import sys, os, time, numpy as np
sys.path.insert(0, os.getcwd())
from oni_ml_b200 import _lib, lda
from oni_ml_b200.corpus import Corpus
rng = np.random.default_rng(0)
V, K, N = 50000, 20, 8000
word = rng.integers(0, V, size=N).astype(np.int32); count = rng.geometric(0.25, size=N).astype(np.int32)
c = Corpus(np.array([0, N], np.int64), word, count, V)
st = _lib.Settings(20, 1e-6, 20, 1e-6, 1)
for g in ("32", "16", "8", "0"):
    if g == "0":
        os.environ["ONI_LDA_LONG_DOC"] = "0"
    else:
        os.environ["ONI_LDA_LONG_DOC"] = "1024"; os.environ["ONI_LDA_LONG_G"] = g
    with lda.LdaEngine(0, K, V, st) as eng:
        eng.set_corpus(c); eng.init_random(1, 2.5)
        eng.em_iteration()
        res = []
        for _ in range(3):
            eng.em_iteration(); s = eng.stats(); res.append((s["ms_estep_kernel"], s["word_visits"]))
        ms, wv = res[-1]
        print("long G=%s: E-step %.2f ms for %d visits -> %.3f us/visit (%d cycles at 1.9 GHz)" % (g, ms, wv, ms * 1e3 / wv, ms * 1e3 / wv * 1900))
