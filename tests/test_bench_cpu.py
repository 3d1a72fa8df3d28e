"""bench.py's bookkeeping that needs no GPU: the ncu records under profiles/ that its roofline
figures come from must describe the kernel sources as they are now (bench.py silently reports
`traffic: null` for a stale record), and the two places that hash those sources must agree."""
import importlib.util
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def test_profile_records_describe_the_current_kernels():
    bench = _load("bench_mod", os.path.join(ROOT, "bench.py"))
    summ = _load("summarize_ncu_mod", os.path.join(ROOT, "profiles", "summarize_ncu.py"))
    assert bench.kernel_source_hash() == summ.source_hash()
    score_sources = ("scoring.cu", "radix_sort.cuh")
    assert bench.kernel_source_hash(score_sources) == summ.source_hash(score_sources)
    est = bench.ncu_profile_record()
    assert est is not None, "profiles/estep_ncu.json is missing or was taken from other kernel sources"
    assert est["dram_bytes_per_visit"] > 100 and 800 < est["fp64_lane_ops_per_visit"] < 1100
    sc = bench.ncu_profile_record("score_ncu.json", score_sources)
    assert sc is not None, "profiles/score_ncu.json is missing or was taken from other kernel sources"
    assert 200 < sc["dram_bytes_per_event"] < 400
    for name in ("estep_ncu.json", "score_ncu.json"):
        rec = json.load(open(os.path.join(ROOT, "profiles", name)))
        assert rec["report"].endswith(".ncu-rep") and rec["duration_ms"] > 0


def test_strong_scaling_cut_is_the_engines_cut():
    """bench.py cuts ONE corpus over the ranks with oni_ml_b200.dist.shard_bounds: contiguous,
    covering, balanced by pairs (the rule of shard_docs in csrc/engine.cu)."""
    import numpy as np
    from oni_ml_b200 import dist as odist
    rng = np.random.default_rng(1)
    lens = rng.integers(0, 200, 5000)
    ptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    for n in (1, 2, 4, 8):
        cut = odist.shard_bounds(ptr, n)
        assert cut[0] == 0 and cut[-1] == 5000 and all(a <= b for a, b in zip(cut, cut[1:]))
        pairs = [int(ptr[b] - ptr[a]) for a, b in zip(cut, cut[1:])]
        assert sum(pairs) == int(ptr[-1])
        assert max(pairs) - min(pairs) <= 2 * int(lens.max())
