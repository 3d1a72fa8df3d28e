"""The branch-free fp64 exp / log / digamma of csrc/devmath.cuh, compiled for the
host (same tables and coefficients as on the device; only the reciprocal seed is a
plain division here) and compared with long-double evaluations.  Bounds: exp within
1.5 ulp; log and the reference's digamma within 4e-15 relative -- the E-step parity
tests on the GPU then bound the device build end to end."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_devmath_accuracy_on_host(tmp_path):
    exe = str(tmp_path / "devmath_host")
    subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-o", exe,
                    os.path.join(ROOT, "tests", "devmath_host.cpp"), "-lm"], check=True)
    out = subprocess.run([exe], check=True, capture_output=True, text=True).stdout
    vals = dict(line.split() for line in out.strip().splitlines())
    assert float(vals["exp_max_ulp_beyond_model"]) < 2.6  # ulp in excess of 0.45 ulp per unit of |x|
    assert float(vals["log_max_rel"]) < 2e-14  # relative to max(|log x|, 1e-3)
    assert float(vals["digamma_max_rel"]) < 4e-15
    assert float(vals["lgamma_max_rel"]) < 4e-15
    assert vals["edge_ok"] == "1"
    # psi_table.h (long-document kernel): the tabulated reference digamma and the short-chain exp
    assert float(vals["psi_table_max_rel"]) < 1e-14  # relative to max(1, |psi|)
    assert float(vals["expk_max_ulp_beyond_model"]) < 3.0  # ulp in excess of 0.6 ulp per unit of |x|
    assert vals["psi_index_ok"] == "1" and vals["expk_edge_ok"] == "1"


def test_generated_tables_are_current(tmp_path):
    """math_tables.h is what gen_math_tables.py writes today."""
    import shutil
    src = os.path.join(ROOT, "oni_ml_b200", "csrc")
    shutil.copy(os.path.join(src, "gen_math_tables.py"), tmp_path / "gen_math_tables.py")
    subprocess.run(["python", "gen_math_tables.py"], cwd=tmp_path, check=True, stdout=subprocess.DEVNULL)
    assert open(tmp_path / "math_tables.h").read() == open(os.path.join(src, "math_tables.h")).read()
