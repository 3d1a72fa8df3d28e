import sys, time, os, numpy as np, subprocess
sys.path.insert(0, os.getcwd())
import torch
from harness.synth import synth_corpus
from oni_ml_b200 import LDA_BIN
ptr, word, count = synth_corpus(1_000_000, 50_000, mu=3.5, seed=20160123, device="cuda", as_torch=True)
ptr, word, count = ptr.cpu().numpy(), word.cpu().numpy(), count.cpu().numpy()
from oni_ml_b200 import lda
from oni_ml_b200.corpus import Corpus
t = time.time()
os.makedirs("/tmp/day", exist_ok=True)
lda.write_corpus("/tmp/day/model.dat", Corpus(ptr, word, count, 50_000))
print("model.dat: %.0f MB written in %.1f s" % (os.path.getsize("/tmp/day/model.dat") / 1e6, time.time() - t), flush=True)
open("/tmp/day/settings.txt", "w").write("var max iter 20\nvar convergence 1e-6\nem max iter 20\nem convergence 1e-6\nalpha estimate\n")
t = time.time()
p = subprocess.run([LDA_BIN, "est", "2.5", "20", "/tmp/day/settings.txt", "1", "/tmp/day/model.dat", "random", "/tmp/day/out"],
                   env=dict(os.environ, ONI_LDA_SEED="4357", ONI_LDA_TIMING="1"), capture_output=True, text=True)
wall = time.time() - t
print("lda est wall %.2f s rc=%d" % (wall, p.returncode))
print(p.stdout[-600:]); print(p.stderr[-900:])
for n in ("likelihood.dat", "final.gamma", "final.beta", "final.other"):
    print(n, os.path.getsize("/tmp/day/out/" + n))
print(open("/tmp/day/out/likelihood.dat").read())
