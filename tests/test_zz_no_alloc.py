"""oni_score_flow_noalloc: the flow scorer for hosts that cannot allocate.

Some callers get a fixed share of the GPU from a pool at start-up (a Spark
executor under a memory arena, a serving process that reserved everything
before taking traffic) or run latency-critical streams next to us; a
cudaMalloc inside the library may fail for them or stall the device.  For those
the library keeps the filter's and the sort's working memory in a fixed array of
its own, good for ONI_NOALLOC_MAX_SURVIVORS (2^18) events below the threshold,
and takes every input and output as caller-owned device memory.

The test feeds the same tables and events to the ordinary entry point
(oni_score_flow, host arrays) and to the allocation-free one (device arrays held
by torch) and requires the identical ranked list -- same event indices, same key
bits, the tolerance the other scoring tests use -- on two calls in a row, the
second with a different threshold so that stale contents of the shared array
would show.  It also checks that too many survivors are refused with a message.

Collected last on purpose (file name), so a run that stops at the first failure
has already been through every other test.
"""
import ctypes as C

import numpy as np
import pytest

from oni_ml_b200 import _lib, scoring

pytestmark = pytest.mark.gpu


def _case(n, D, V, K, seed):
    rng = np.random.default_rng(seed)
    theta = np.ascontiguousarray(rng.dirichlet(np.full(K, 0.3), size=D))
    pwz = np.ascontiguousarray(rng.dirichlet(np.full(V, 0.05), size=K).T)
    ev = []
    for lim in (D, D, V, V):
        a = rng.integers(0, lim, size=n).astype(np.int32)
        a[rng.random(n) < 0.01] = -1  # lookup misses -> the 0.05 vector
        ev.append(a)
    return theta, pwz, ev


def _noalloc(theta_d, pwz_d, ev_d, K, tol, cap, out_idx, out_key):
    """One call of the allocation-free entry point through the C ABI."""
    L = _lib.load()
    n_out = C.c_int64(-1)
    rc = L.oni_score_flow_noalloc(0, K, theta_d.data_ptr(), pwz_d.data_ptr(), ev_d[0].numel(),
                                  ev_d[0].data_ptr(), ev_d[1].data_ptr(), ev_d[2].data_ptr(),
                                  ev_d[3].data_ptr(), float(tol), cap, out_idx.data_ptr(),
                                  out_key.data_ptr(), C.byref(n_out), None)
    return rc, n_out.value


def test_noalloc_matches_allocating_entry_point_twice_in_a_row():
    import torch
    n, D, V, K = 300000, 4000, 2500, 20
    theta, pwz, ev = _case(n, D, V, K, 11)
    dev = torch.device("cuda", 0)
    theta_d, pwz_d = torch.from_numpy(theta).to(dev), torch.from_numpy(pwz).to(dev)
    ev_d = [torch.from_numpy(a).to(dev) for a in ev]
    everything = scoring.score_flow(theta, pwz, *ev, np.inf, sides=False)
    # the caller owns the outputs as well; allocate them once, before the calls
    cap = 1 << 16
    out_idx = torch.full((cap,), -7, dtype=torch.int64, device=dev)
    out_key = torch.full((cap,), -7.0, dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    for call, q in enumerate((0.05, 0.002)):  # two calls in a row; the second keeps fewer events
        tol = float(everything["key"][int(q * n)])
        want = scoring.score_flow(theta, pwz, *ev, tol, out_cap=cap, sides=False)
        before = torch.cuda.memory_allocated(dev)
        free_before = torch.cuda.mem_get_info(dev)[0]
        rc, got_n = _noalloc(theta_d, pwz_d, ev_d, K, tol, cap, out_idx, out_key)
        assert rc == 0, _lib.load().oni_last_error()
        assert torch.cuda.memory_allocated(dev) == before
        if call == 1:  # (the first call may make the driver load a kernel image)
            assert torch.cuda.mem_get_info(dev)[0] == free_before
        m = min(cap, want["n"])
        assert got_n == want["n"] and m > 100
        assert np.array_equal(out_idx[:m].cpu().numpy(), want["idx"])
        assert np.array_equal(out_key[:m].cpu().numpy().view(np.int64), want["key"].view(np.int64))


def test_noalloc_refuses_more_survivors_than_its_scratch_holds():
    import torch
    n, D, V, K = (1 << 18) + 5000, 300, 200, 20
    theta, pwz, ev = _case(n, D, V, K, 12)
    dev = torch.device("cuda", 0)
    theta_d, pwz_d = torch.from_numpy(theta).to(dev), torch.from_numpy(pwz).to(dev)
    ev_d = [torch.from_numpy(a).to(dev) for a in ev]
    out_idx = torch.zeros(16, dtype=torch.int64, device=dev)
    out_key = torch.zeros(16, dtype=torch.float64, device=dev)
    rc, _ = _noalloc(theta_d, pwz_d, ev_d, K, 10.0, 16, out_idx, out_key)  # every event survives
    assert rc == 1 and b"fixed scratch" in _lib.load().oni_last_error()
    rc, _ = _noalloc(theta_d, pwz_d, ev_d, K, 10.0, (1 << 18) + 1, out_idx, out_key)
    assert rc == 1 and b"out_cap" in _lib.load().oni_last_error()
    # and the array is still usable afterwards
    rc, got_n = _noalloc(theta_d, pwz_d, ev_d, K, 1e-20, 16, out_idx, out_key)
    assert rc == 0 and got_n == 0
