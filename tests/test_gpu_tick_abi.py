"""The closed control tick through the C ABI ONLY (qrw_tick_*, include/qrw_b200.h): ctypes + numpy, no torch, no Python
composition -- what a C++ caller of scripts/Controller.py:222-303's replacement reaches.  Tick by tick against the chain
composed from the oracle's pieces, then whole MPC periods (one CUDA-graph launch each, host buffers in and out) against the
tick-by-tick path of a second handle."""
import ctypes as C

import numpy as np
import pytest

from conftest import assert_forces_close
from test_gpu_controller import H_REF, _oracle_chain

pytestmark = pytest.mark.gpu


def _lib():
    from qrw_b200._lib import load_library

    return load_library()


def _create(L, B):
    h = C.c_void_p()
    assert L.qrw_tick_create(C.byref(h), B, 0.002, 0.02, 0.32, 0.32, 20, H_REF, None, 0) == 0, L.qrw_last_error()
    return h


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def test_tick_by_tick_from_host_pointers_matches_the_oracle_chain(oracle):
    from qrw_b200 import workloads as wl

    L, B, T = _lib(), 32, 22
    W = wl.PlannerWorkload(B, 31, change_every=1000, yaw0=0.0, speed=0.5)
    h = _create(L, B)
    ref_tick = _oracle_chain(oracle, B)
    tau, qd, vd, fr = (np.empty((B, 12)) for _ in range(4))
    for k in range(T):
        q7, hv, vr, code = W.tick(k)
        assert L.qrw_tick_run_host(h, k, _p(q7), _p(hv), _p(vr), _p(code)) == 0, L.qrw_last_error()
        assert L.qrw_tick_result_host(h, _p(tau), _p(qd), _p(vd), _p(fr)) == 0
        ref, _ = ref_tick(k, q7, hv, vr, code)
        assert_forces_close(fr, ref["f_res"], f"f_res tick {k}")
        np.testing.assert_allclose(qd, ref["qdes"], rtol=1e-8, atol=1e-8, err_msg=f"qdes tick {k}")
        np.testing.assert_allclose(vd, ref["vdes"], rtol=1e-7, atol=1e-7, err_msg=f"vdes tick {k}")
        np.testing.assert_allclose(tau, ref["tau_ff"], rtol=1e-5, atol=2e-4, err_msg=f"tau tick {k}")
    assert L.qrw_tick_destroy(h) == 0


def test_period_host_equals_tick_by_tick_and_refuses_before_capture():
    from qrw_b200 import workloads as wl

    L, B, K = _lib(), 200, 10
    W = wl.PlannerWorkload(B, 32, change_every=1000, yaw0=0.0, speed=0.5)
    seq = [W.tick(k) for k in range(3 * K)]
    a, g = _create(L, B), _create(L, B)
    assert L.qrw_tick_capture_period(g) != 0 and b"first tick" in L.qrw_last_error()
    assert L.qrw_tick_period_host(g, _p(seq[0][0]), _p(seq[0][1]), _p(seq[0][2]), None, None, None, None) != 0
    for k in range(K):
        for h in (a, g):
            assert L.qrw_tick_run_host(h, k, *[_p(x) for x in seq[k]]) == 0, L.qrw_last_error()
    assert L.qrw_tick_capture_period(g) == 0, L.qrw_last_error()
    tau = np.empty((B, 12))
    out = [np.empty((K, B, 12)) for _ in range(3)]
    for period in (1, 2):
        ticks = seq[K * period:K * (period + 1)]
        want = []
        for i, (q7, hv, vr, code) in enumerate(ticks):
            assert L.qrw_tick_run_host(a, K * period + i, _p(q7), _p(hv), _p(vr), _p(code)) == 0
            assert L.qrw_tick_result_host(a, _p(tau), None, None, None) == 0
            want.append(tau.copy())
        stk = [np.ascontiguousarray(np.stack([t[j] for t in ticks])) for j in range(4)]
        assert L.qrw_tick_period_host(g, _p(stk[0]), _p(stk[1]), _p(stk[2]), _p(stk[3]), *[_p(o) for o in out]) == 0, L.qrw_last_error()
        for i in range(K):
            assert np.array_equal(out[0][i], want[i]), (period, i)
    for h in (a, g):
        assert L.qrw_tick_destroy(h) == 0
