"""Behaviour on inputs the reference never sees in its own tests: a robot that attaches to the MPC server
mid-run, and non-finite inputs.  The MPC QP itself is always feasible for finite inputs (states are free, f = 0 satisfies the
cone and force-enable rows), so OSQP's infeasibility certificates (eps_prim_inf / eps_dual_inf, src/MPC.cpp:528-529) can only
fire on garbage input; the kernels do not implement them and this file pins what they do instead."""
import numpy as np
import pytest

from conftest import assert_forces_close

pytestmark = pytest.mark.gpu


def test_a_slots_first_request_is_a_cold_start_whatever_index_it_carries(oracle):
    from qrw_b200 import workloads as wl
    from qrw_b200.mpc_server import MPCServer, SharedBlock, compress_data_in

    blk = SharedBlock(robots=2, create=True)
    srv = MPCServer(blk)
    try:
        w = wl.MpcWorkload(1, 77, gaits=("trot",), ticks=3)
        xr, fs = w.inputs(0)
        xref, fsteps = xr.reshape(17, 12).T, fs.reshape(12, 20).T
        # robot 1 attaches with k = 70 (MPC call index 7): without the per-slot flag the update path would run on a zero state
        blk.data_in[1, :] = compress_data_in(70, 10, xref, fsteps)
        blk.new_data[1] = 1
        assert srv.poll() == 1 and blk.new_result[1] == 1
        got = blk.data_out[1].reshape(16, 24).T
        o = oracle.OracleMPC(1)
        o.run(0, xr, fs)
        assert np.isfinite(got).all()
        assert_forces_close(got, o.get_latest_result().reshape(16, 24).T, "first request served as a cold start")
        # its second request takes the update path as usual
        x1 = o.get_latest_result()[:, :12].copy()
        xr, fs = w.inputs(1, x1)
        blk.new_result[1] = 0
        blk.data_in[1, :] = compress_data_in(80, 10, xr.reshape(17, 12).T, fs.reshape(12, 20).T)
        blk.new_data[1] = 1
        assert srv.poll() == 1
        o.run(1, xr, fs)
        assert_forces_close(blk.data_out[1].reshape(16, 24).T, o.get_latest_result().reshape(16, 24).T, "second request: update path")
    finally:
        srv.close()
        blk.close(unlink=True)


def test_non_finite_input_of_one_robot_does_not_touch_its_neighbours(oracle):
    from qrw_b200 import BatchedMPC, workloads as wl

    B = 8
    w = wl.MpcWorkload(B, 78, gaits=("trot", "walk"), ticks=2)
    xr, fs = w.inputs(0)
    bad = xr.copy()
    bad[3, 2 + 12 * 4] = np.nan  # robot 3: NaN height reference at stage 4
    gpu, cpu = BatchedMPC(B), oracle.OracleMPC(B)
    gpu.set_max_iter(100)
    cpu.set_max_iter(100)
    gpu.run(0, bad, fs)
    cpu.run(0, xr, fs, 4)
    rg, rc = gpu.get_latest_result(), cpu.get_latest_result()
    ok = np.arange(B) != 3
    assert np.array_equal(gpu.info()["iters"][ok], cpu.info()["iters"][ok])
    assert_forces_close(rg[ok], rc[ok], "robots next to the one with NaN input")
    # the robot itself behaves as OSQP v0.6 does on NaN data (the oracle restates it): the inf-norms of the residuals skip NaN
    # (`if (abs_v_i > max)`), so the first termination check reports "solved" with a NaN solution.  The reference never reads
    # the status (MPC.cpp:569-599 copies the solution out); what a caller can rely on is that the forces are not finite.
    cpu.run(0, bad, fs, 4)
    assert not np.isfinite(rg[3]).all()
    assert gpu.info()["status"][3] == cpu.info()["status"][3]
    assert gpu.info()["iters"][3] == cpu.info()["iters"][3]


def test_pinned_host_buffers_take_the_zero_copy_path_with_identical_results():
    """qrw_mpc_run_host with PINNED inputs: the kernel reads xref / fsteps straight from host memory and mirrors every result into
    the handle's pinned result buffer (no staging copies).  Same bits as the pageable path: results (read through the pinned
    buffer and through an ordinary array), iteration counts, gait, three ticks of a mixed-gait closed loop; and a device-side run
    afterwards must not be answered from the stale host mirror."""
    import torch

    from qrw_b200 import BatchedMPC, workloads as wl

    B = 600  # more robots than CTA slots: CTAs take several robots each
    w = wl.MpcWorkload(B, 23, gaits=("trot", "walk", "bounding", "pacing", "static"), ticks=4)
    a, b = BatchedMPC(B), BatchedMPC(B)
    px, pf, pr = b.pinned()
    x0 = None
    for t in range(3):
        xr, fs = w.inputs(t, x0)
        a.run(t, xr, fs)  # pageable numpy arrays: staged
        px[:], pf[:] = xr, fs
        pr[:] = np.nan
        b.run(t, px, pf)  # the handle's pinned buffers: zero copy
        ra = a.get_latest_result()
        rb_pinned = b.get_latest_result(out=pr)
        assert rb_pinned is pr and np.array_equal(ra, pr), f"tick {t}: pinned result buffer"
        assert np.array_equal(ra, b.get_latest_result()), f"tick {t}: ordinary result array"
        assert np.array_equal(a.info()["iters"], b.info()["iters"]) and np.array_equal(a.get_gait(), b.get_gait())
        x0 = ra[:, :12].copy()
    # a torch pinned tensor is the same kind of memory
    xr, fs = w.inputs(3, x0)
    tx, tf = torch.from_numpy(xr).pin_memory(), torch.from_numpy(fs).pin_memory()
    a.run(3, xr, fs)
    b.run(3, tx.numpy(), tf.numpy())
    assert np.array_equal(a.get_latest_result(), b.get_latest_result(out=pr))
    # device-side run: the mirror is stale now, the getter must copy
    dx, df = torch.from_numpy(w.inputs(4 if False else 3, x0)[0]).cuda(), torch.from_numpy(fs).cuda()
    pr[:] = np.nan
    b.run(4, dx, df)
    a.run(4, xr, fs)
    torch.cuda.synchronize()
    assert np.array_equal(a.get_latest_result(), b.get_latest_result(out=pr))
