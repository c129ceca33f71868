"""MPC server transport: the reference's MPC_Wrapper shared-memory wire format
(scripts/MPC_Wrapper.py:47-52,227-264) for many robot processes on one GPU.  CPU: wire format, shared block across
processes, client-side result buffering.  GPU: robot processes posting at different times are served by the indexed MPC
kernel and each receives what the oracle computes for ITS call history."""
import multiprocessing as mp
import time

import numpy as np
import pytest

from qrw_b200 import workloads as wl
from qrw_b200.mpc_server import MPCClient, SharedBlock, compress_data_in, slot_sizes

K_MPC = 10
Q_INIT = np.array([0.0, 0.0, 0.2447495, 0.0, 0.0, 0.0, 1.0] + list(wl.Q_INIT))


def _robot_inputs(seed, T):
    """(xref 12x17, fsteps 20x12, gait 20x4) of T consecutive MPC calls of one robot (open loop: fixed initial state)."""
    w = wl.MpcWorkload(1, seed, gaits=("trot", "walk"), ticks=T + 1)
    out = []
    for t in range(T):
        xr, fs = w.inputs(t)
        out.append((xr.reshape(17, 12).T.copy(), fs.reshape(12, 20).T.copy(), w.gait(t)[0]))
    return out


def test_wire_format_is_the_reference_one():
    xref, fsteps, _ = _robot_inputs(3, 1)[0]
    fs = fsteps.copy()
    fs[19, 5] = np.nan
    d = compress_data_in(30, K_MPC, xref, fs)
    per_in, per_out = slot_sizes()
    assert d.shape == (per_in,) == (1 + 17 * 12 + 12 * 20,) and per_out == 24 * 16
    assert d[0] == 3.0 and (d[1:205].reshape(12, 17) == xref).all() and (d[205:].reshape(20, 12)[:19] == fsteps[:19]).all()
    assert d[205 + 19 * 12 + 5] == 0.0  # NaN footsteps are zeroed (MPC_Wrapper.py:236)


def _child_post(name, slot):
    blk = SharedBlock(name=name, robots=3)
    c = MPCClient(blk, slot, K_MPC, Q_INIT)
    xref, fsteps, gait = _robot_inputs(10 + slot, 1)[0]
    c.solve(0, xref, fsteps, gait)
    blk.close()


def test_shared_block_across_processes_and_client_buffering():
    blk = SharedBlock(robots=3, create=True)
    try:
        ctx = mp.get_context("spawn")
        ps = [ctx.Process(target=_child_post, args=(blk.name, s)) for s in (0, 2)]
        [p.start() for p in ps]
        [p.join(60) for p in ps]
        assert list(blk.new_data) == [1, 0, 1] and blk.running[0] == 1
        xref, _, _ = _robot_inputs(12, 1)[0]
        assert (blk.data_in[2, 1:205].reshape(12, 17) == xref).all() and blk.data_in[1].sum() == 0
        # client-side semantics of MPC_Wrapper.get_latest_result (:104-126): default block first, then the last result
        c = MPCClient(blk, 1, K_MPC, Q_INIT)
        first = c.get_latest_result()
        assert (first[12:, 0] == np.array([0.0, 0.0, 8.0] * 4)).all() and first[2, 0] == Q_INIT[2]
        assert c.get_latest_result() is c.last_available_result  # nothing new: the old result again
        res = np.arange(384.0)
        blk.data_out[1, :] = res
        blk.new_result[1] = 1
        got = c.get_latest_result()
        assert (got == res.reshape(24, 16, order="F")).all() and blk.new_result[1] == 0
        c.stop_parallel_loop()
        assert blk.running[0] == 0
    finally:
        blk.close(unlink=True)


def _robot_process(name, slot, seed, T, delay, q):
    """A robot's control process: numpy only.  Posts T MPC requests, each after the previous result has arrived."""
    blk = SharedBlock(name=name, robots=4)
    c = MPCClient(blk, slot, K_MPC, Q_INIT)
    c.get_latest_result()  # the default block of the first control tick
    time.sleep(delay)
    got = []
    for t, (xref, fsteps, gait) in enumerate(_robot_inputs(seed, T)):
        c.solve(t * K_MPC, xref, fsteps, gait)
        t0 = time.time()
        while not blk.new_result[slot]:
            if time.time() - t0 > 120:
                q.put((slot, None))
                return
            time.sleep(0.0005)
        got.append(c.get_latest_result().copy())
    q.put((slot, np.stack(got)))
    blk.close()


@pytest.mark.gpu
def test_server_serves_staggered_robot_processes(oracle):
    from conftest import assert_forces_close
    from qrw_b200.mpc_server import MPCServer

    plan = {0: (21, 4, 0.0), 1: (22, 2, 0.15), 3: (23, 5, 0.05)}  # slot -> (seed, calls, start delay); slot 2 never posts
    blk = SharedBlock(robots=4, create=True)
    srv = MPCServer(blk)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_robot_process, args=(blk.name, s, seed, T, d, q)) for s, (seed, T, d) in plan.items()]
    try:
        [p.start() for p in ps]
        res, served, t0 = {}, 0, time.time()
        while len(res) < len(plan) and time.time() - t0 < 300:
            served += srv.poll()
            while not q.empty():
                s, r = q.get()
                res[s] = r
            time.sleep(0.0002)
        [p.join(30) for p in ps]
        assert served == sum(T for _, T, _ in plan.values())
        for s, (seed, T, _) in plan.items():
            assert res[s] is not None and res[s].shape == (T, 24, 16)
            o = oracle.OracleMPC(1)
            for t, (xref, fsteps, _) in enumerate(_robot_inputs(seed, T)):
                o.run(t, xref.ravel(order="F")[None, :], fsteps.ravel(order="F")[None, :])
                want = o.get_latest_result().reshape(16, 24).T
                assert_forces_close(res[s][t], want, f"slot {s} call {t}")
        assert (blk.data_out[2] == 0).all() and blk.new_result[2] == 0  # the silent slot was never touched
        # the native loop: stops when `running` is cleared
        blk.running[0] = 0
        assert srv.serve() == 0
    finally:
        [p.kill() for p in ps if p.is_alive()]
        srv.close()
        blk.close(unlink=True)
