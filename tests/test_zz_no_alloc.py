"""qrw_wbctick_noalloc_device (csrc/noalloc_abi.cuh): the whole-body tick with its scratch in a fixed file-scope __device__ array,
for callers that cannot call cudaMalloc.  No counterpart in the reference.  Called
through the C ABI (ctypes) and compared with the existing entry point -- the first qrw_wbctick_run_device of a fresh handle, which
is the same cold-started tick -- on two calls in a row (the second call finds the first call's contents in the shared scratch),
then with the oracle.  Named test_zz_* so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

from conftest import assert_forces_close

pytestmark = pytest.mark.gpu


def _noalloc_tick(L, B, args, stream=None):
    import torch

    dev = torch.device("cuda", 0)
    tin = [torch.from_numpy(np.ascontiguousarray(a, np.float64)).to(dev) for a in args]
    outs = [torch.full((B, 12), np.nan, dtype=torch.float64, device=dev) for _ in range(4)]  # tau_ff, qdes, vdes, f_res
    torch.cuda.synchronize()
    rc = L.qrw_wbctick_noalloc_device(B, 0, *[C.c_void_p(t.data_ptr()) for t in tin], *[C.c_void_p(t.data_ptr()) for t in outs],
                                      C.c_void_p(stream or 0))
    torch.cuda.synchronize()
    return rc, [o.cpu().numpy() for o in outs]


def _handle_tick(B, args):
    from qrw_b200 import BatchedWbcTick

    t = BatchedWbcTick(B)  # a fresh handle: its first tick is the cold start the no-alloc entry point always does
    t.compute(*args)
    out = [t.tau_ff.copy(), t.qdes.copy(), t.vdes.copy(), t.f_with_delta.copy()]
    t.close()
    return out


def test_no_alloc_tick_equals_the_handle_entry_point_on_two_calls_in_a_row(oracle):
    from qrw_b200 import _lib, workloads as wl

    L = _lib.load_library()
    # two calls in a row with different sizes and inputs: the second one finds the first one's leftovers in the scratch
    for call, (B, seed, level) in enumerate(((1000, 71, True), (4096, 72, False), (37, 73, False))):
        args = wl.wbc_tick_inputs(B, seed=seed, level_base=level)
        rc, got = _noalloc_tick(L, B, args)
        assert rc == 0, L.qrw_last_error().decode()
        want = _handle_tick(B, args)
        for name, g, w in zip(("tau_ff", "qdes", "vdes"), got[:3], want[:3]):
            assert np.isfinite(g).all(), f"call {call}: {name} not written"
            np.testing.assert_allclose(g, w, rtol=1e-10, atol=1e-10, err_msg=f"call {call}: {name}")
        assert_forces_close(got[3], want[3], f"call {call}: f_res")
        if B <= 1000:  # and against the oracle's composition of wbc_controller.compute (cold start)
            ref = oracle.wbc_tick(oracle.OracleWBC(B), *args, nthreads=8)
            assert_forces_close(got[3], ref["f_res"], f"call {call}: f_res vs oracle")
            np.testing.assert_allclose(got[0], ref["tau_ff"], rtol=1e-5, atol=1e-4)
            np.testing.assert_allclose(got[1], ref["qdes"], rtol=1e-10, atol=1e-10)


def test_no_alloc_tick_rejects_what_does_not_fit():
    from qrw_b200 import _lib, workloads as wl

    L = _lib.load_library()
    B = 4097  # QRW_WBCTICK_NOALLOC_MAX_ROBOTS + 1
    rc, got = _noalloc_tick(L, B, wl.wbc_tick_inputs(B, seed=74))
    assert rc != 0 and "4096" in L.qrw_last_error().decode() and "exceeds" in L.qrw_last_error().decode()
    assert np.isnan(got[0]).all()  # nothing was launched
    assert L.qrw_wbctick_noalloc_device(0, 0, *[C.c_void_p(8)] * 11, C.c_void_p(0)) != 0
    assert L.qrw_wbctick_noalloc_device(16, 0, *[C.c_void_p(0)] * 11, C.c_void_p(0)) != 0
