import sys, os, numpy as np
sys.path.insert(0, os.getcwd())
from oni_ml_b200 import _lib, lda
from oni_ml_b200.corpus import Corpus
rng = np.random.default_rng(0)
V, K, N = 50000, 20, 8000
word = rng.integers(0, V, size=N).astype(np.int32); count = rng.geometric(0.25, size=N).astype(np.int32)
c = Corpus(np.array([0, N], np.int64), word, count, V)
st = _lib.Settings(20, 1e-6, 20, 1e-6, 1)
with lda.LdaEngine(0, K, V, st) as eng:
    eng.set_corpus(c); eng.init_random(1, 2.5)
    for _ in range(3):
        eng.em_iteration()
    print(eng.stats())
