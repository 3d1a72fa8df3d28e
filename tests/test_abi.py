"""The C-ABI library loads, exports every symbol include/oni_lda.h declares, and
its host-only entry points (settings / corpus parser / writers) honour the
reference's file grammar byte for byte.  No GPU needed; no compute calls."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from oni_ml_b200 import _lib, files, lda

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "oni_lda.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(oni_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = C.CDLL(_lib.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 40
    for n in names:
        assert hasattr(lib, n), "liboni_lda.so lacks %s" % n
    # and the binding table covers the header
    assert set(names) == set(_lib.SIGNATURES), set(names) ^ set(_lib.SIGNATURES)
    assert _lib.load().oni_version().startswith(b"oni-lda-b200")


def test_read_settings_shipped_and_fixed(tmp_path, golden_dir):
    st = lda.read_settings(os.path.join(golden_dir, "c0_k20", "settings.txt"))
    assert (st.var_max_iter, st.em_max_iter, st.estimate_alpha) == (20, 20, 1)
    # thresholds are floats (lda-inference.h:10): 1e-6 -> 9.99999997e-07
    assert float(st.var_converged) == float(np.float32(1e-6)) != 1e-6
    st = lda.read_settings(os.path.join(golden_dir, "c0_k5_fixed", "settings.txt"))
    assert (st.var_max_iter, st.em_max_iter, st.estimate_alpha) == (5, 6, 0)
    st = lda.read_settings(os.path.join(golden_dir, "c0_k37_unbounded", "settings.txt"))
    assert st.var_max_iter == -1
    with pytest.raises(_lib.OniError) as e:
        lda.read_settings(str(tmp_path / "nope.txt"))
    assert e.value.code == 2


def test_read_corpus_matches_python_reader(golden_dir, c0_corpus):
    c = lda.read_corpus(os.path.join(golden_dir, "c0_model.dat"))
    assert c.num_docs == c0_corpus.num_docs == 403
    assert c.num_terms == c0_corpus.num_terms == 300
    assert np.array_equal(c.doc_ptr, c0_corpus.doc_ptr)
    assert np.array_equal(c.word, c0_corpus.word)
    assert np.array_equal(c.count, c0_corpus.count)
    # the grammar's edge cases survive: empty document, repeated word id
    lens = np.diff(c.doc_ptr)
    assert lens[-3] == 0 and lens[-1] == 3 and list(c.word[-3:]) == [11, 11, 11]


def test_read_corpus_rejects_garbage(tmp_path):
    p = tmp_path / "bad.dat"
    p.write_text("2 1:1 x:3\n")
    with pytest.raises(_lib.OniError) as e:
        lda.read_corpus(str(p))
    assert e.value.code == 2
    with pytest.raises(_lib.OniError):
        lda.read_corpus(str(tmp_path / "missing.dat"))


def test_writers_byte_identical_to_reference(tmp_path, golden_dir):
    g = os.path.join(golden_dir, "c0_k20")
    beta = files.read_matrix(os.path.join(g, "final.beta"))
    other = files.read_other(os.path.join(g, "final.other"))
    gamma = files.read_matrix(os.path.join(g, "final.gamma"))
    lda.save_model(str(tmp_path / "final"), beta, other["alpha"])
    lda.save_gamma(str(tmp_path / "final.gamma"), gamma)
    for name in ("final.beta", "final.other", "final.gamma"):
        assert open(tmp_path / name).read() == open(os.path.join(g, name)).read(), name


def test_compute_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(_lib.OniError) as e:
        lda.LdaEngine(0, 20, 300, _lib.Settings.shipped())
    assert e.value.code == 3 and "no CPU path" in str(e.value)
    th = np.full((2, 20), 0.05)
    from oni_ml_b200 import scoring
    with pytest.raises(_lib.OniError) as e:
        scoring.score_dns(th, th, [0], [0], 1.0)
    assert e.value.code == 3


def test_cli_usage_and_errors(tmp_path):
    p = lda.run_cli([])
    assert "usage : lda est" in p.stdout and p.returncode == 0
    p = lda.run_cli(["est", "2.5", "20", str(tmp_path / "nosettings"), "1", "x", "random", str(tmp_path)],
                    check_rc=False)
    assert p.returncode == 2 and "cannot open settings" in p.stderr


def test_fast_fixed10_formatter_is_exact():
    """The writers print "%5.10f" with an exact 128-bit conversion instead of printf;
    it must give printf's digits on ties, near-ties, tiny, negative and large values."""
    import random
    import struct
    L = C.CDLL(_lib.LIB_PATH)
    L.oni_internal_format_fixed10.argtypes = [C.c_double, C.c_char_p]
    L.oni_internal_format_fixed10.restype = C.c_int
    buf = C.create_string_buffer(128)
    rnd = random.Random(1)
    vals = [0.0, -0.0, 1.0, -1.0, 0.5e-10, 1.5e-10, 2.5e-10, 1e-11, -1e-11, 123456789.123456789, 1e15,
            9.0e18 * 0.99, 5e-324, -5e-324, 2.5, -100.0, 0.12345678905, 0.12345678915, 1e-300,
            0.99999999995, 0.999999999949999, 9.99999999995]
    for _ in range(60000):
        vals.append(rnd.choice([-1, 1]) * rnd.random() * 2.0 ** rnd.uniform(-40, 40))
    for _ in range(40000):
        v = (rnd.randrange(0, 10 ** 13) * 10 + 5) / 1e11  # ties at the 11th decimal and neighbours
        vals += [v, float(np.nextafter(v, 0)), float(np.nextafter(v, 1e9))]
    for _ in range(40000):
        vals.append(struct.unpack("<d", struct.pack("<Q", rnd.getrandbits(64)))[0])
    checked = 0
    for v in vals:
        n = L.oni_internal_format_fixed10(v, buf)
        if n < 0:
            assert not abs(v) < 9.0e18  # non-finite or huge: the writer uses fprintf for those
            continue
        assert buf.value.decode() == "%5.10f" % v, repr(v)
        checked += 1
    assert checked > 150000


def test_writer_handles_nonfinite_and_huge(tmp_path):
    g = np.array([[1.5, float("inf"), -2.0], [1e300, float("nan"), 0.0]])
    lda.save_gamma(str(tmp_path / "g"), g)
    want = "".join(" ".join("%5.10f" % v for v in row) + "\n" for row in g)
    assert open(tmp_path / "g").read() == want


def test_lda_pre_byte_identical_to_reference_script(tmp_path, golden_dir):
    """oni_lda_pre against what the reference's lda_pre.py wrote for the same doc_wc.dat
    (tests/golden/pre, oracle/make_golden.py): index mappings must be bit-exact."""
    import shutil
    src = os.path.join(golden_dir, "pre")
    shutil.copy(os.path.join(src, "doc_wc.dat"), tmp_path / "doc_wc.dat")
    lda.lda_pre(str(tmp_path) + "/")
    for name in ("words.dat", "doc.dat", "model.dat"):
        assert open(tmp_path / name, "rb").read() == open(os.path.join(src, name), "rb").read(), name
    # and the model.dat it wrote parses back to the same pairs (repeated words kept)
    c = lda.read_corpus(str(tmp_path / "model.dat"))
    assert c.nnz == 1501 and c.num_docs == len(open(tmp_path / "doc.dat").readlines())
    with pytest.raises(_lib.OniError):
        lda.lda_pre(str(tmp_path / "nowhere") + "/")


def test_product_never_touches_the_oracle():
    """The oracle is test infrastructure: nothing under oni_ml_b200/ (Python or C/CUDA) may
    import, include, link or execute anything under oracle/."""
    pkg = os.path.join(ROOT, "oni_ml_b200")
    offenders = []
    for base, _dirs, names in os.walk(pkg):
        if os.sep + "build" in base:
            continue
        for n in names:
            if n.endswith((".py", ".c", ".cu", ".cuh", ".h", ".inc")) or n == "Makefile":
                text = open(os.path.join(base, n), errors="replace").read()
                if re.search(r"(from|import)\s+oracle|oracle/|pyoracle|liboni_oracle|lda_ref", text):
                    offenders.append(os.path.join(base, n))
    assert offenders == []
    # and the shared library has no dependency on it
    import subprocess
    deps = subprocess.run(["ldd", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in deps and "torch" not in deps


def test_product_package_needs_no_torch():
    """north_star: "no PyTorch".  The Python side of the product is a ctypes binding; importing every
    module of oni_ml_b200 must not pull torch in (the corpus generator and the torchrun bootstrap,
    which do use it, live under harness/)."""
    import subprocess
    import sys
    code = ("import sys, oni_ml_b200, oni_ml_b200._lib, oni_ml_b200.lda, oni_ml_b200.scoring, oni_ml_b200.files, "
            "oni_ml_b200.corpus, oni_ml_b200.dist, oni_ml_b200.events; oni_ml_b200._lib.load(); "
            "print('torch' in sys.modules)")
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=ROOT)
    assert out.returncode == 0, out.stderr
    assert out.stdout.strip() == "False"


def test_write_corpus_is_the_inverse_of_read_corpus(tmp_path, c0_corpus, golden_dir):
    """oni_lda_write_corpus writes the model.dat text lda_pre.py:94 writes (the committed fixture
    byte for byte) and what it writes parses back to the same arrays."""
    p = str(tmp_path / "model.dat")
    lda.write_corpus(p, c0_corpus)
    assert open(p).read() == open(os.path.join(golden_dir, "c0_model.dat")).read()
    back = lda.read_corpus(p)
    assert np.array_equal(back.doc_ptr, c0_corpus.doc_ptr) and np.array_equal(back.word, c0_corpus.word)
    assert np.array_equal(back.count, c0_corpus.count)
    with pytest.raises(_lib.OniError):
        lda.write_corpus(str(tmp_path / "no" / "dir" / "model.dat"), c0_corpus)
