"""The closed control tick in HBM (qrw_b200.controller.BatchedController = scripts/Controller.py:222-303 for B robots:
planners -> MPC every k_mpc-th tick -> whole-body tick) against the same chain composed from the oracle's pieces, tick by
tick on the same seeded inputs; then one MPC period replayed from a CUDA graph against the tick-by-tick path."""
import numpy as np
import pytest

from conftest import assert_forces_close

pytestmark = pytest.mark.gpu
H_REF = 0.2447495


def _oracle_chain(oracle, B):
    from qrw_b200 import workloads as wl

    st = dict(P=oracle.OraclePlanner(B, h_ref=H_REF), M=oracle.OracleMPC(B), W=oracle.OracleWBC(B), res=None, first=True,
              qdes=np.tile(wl.Q_INIT, (B, 1)), vdes=np.zeros((B, 12)), p=np.zeros((B, 12)), v=np.zeros((B, 12)))

    def tick(k, q7, hv, vr, code):
        st["P"].run(k, q7, hv, vr, code, stages=15)
        o = st["P"].outputs()
        if k % 10 == 0:
            st["M"].run(k, o["xref"], o["fsteps"], 8)
            st["res"] = st["M"].get_latest_result()
        if st["first"]:  # MPC_Wrapper.py:64-71,122-126: default forces on the very first call
            f_cmd, st["first"] = np.tile([0.0, 0.0, 8.0] * 4, (B, 1)), False
        else:
            f_cmd = st["res"][:, 12:24].copy()
        p, v, a = oracle.feet_commands(o["fpos"], o["fvel"], o["facc"], q7, vr, st["p"], st["v"], H_REF)
        st["p"], st["v"] = p, v
        q19, dq18 = np.zeros((B, 19)), np.zeros((B, 18))
        q19[:, 2], q19[:, 6], q19[:, 7:] = H_REF, 1.0, st["qdes"]
        dq18[:, :6], dq18[:, 6:] = vr, st["vdes"]
        contacts = np.ascontiguousarray(o["gait"].reshape(B, 4, 20)[:, :, 0])
        ref = oracle.wbc_tick(st["W"], q19, dq18, f_cmd, contacts, p, v, a, nthreads=8)
        st["qdes"], st["vdes"] = ref["qdes"], ref["vdes"]
        return ref, f_cmd

    return tick


def test_closed_loop_tick_matches_oracle_chain(oracle):
    import torch

    from qrw_b200 import workloads as wl
    from qrw_b200.controller import BatchedController

    B, T = 48, 24
    W = wl.PlannerWorkload(B, 21, change_every=1000, yaw0=0.0, speed=0.5)
    ctl = BatchedController(B, h_ref=H_REF)
    ref_tick = _oracle_chain(oracle, B)
    res = ctl.results()
    for k in range(T):
        q7, hv, vr, code = W.tick(k)
        ctl.tick(k, *[torch.from_numpy(a).cuda() for a in (q7, hv, vr, code)])
        torch.cuda.synchronize()
        ref, f_cmd = ref_tick(k, q7, hv, vr, code)
        if k > 0:  # f_cmd of tick k > 0 is the oracle MPC's first-step force block
            assert_forces_close(ctl.mpc_result.cpu().numpy()[:, 12:24], f_cmd, f"MPC f0 tick {k}")
        assert_forces_close(res["f_res"].cpu().numpy(), ref["f_res"], f"f_res tick {k}")
        np.testing.assert_allclose(res["q_des"].cpu().numpy(), ref["qdes"], rtol=1e-8, atol=1e-8, err_msg=f"qdes tick {k}")
        np.testing.assert_allclose(res["v_des"].cpu().numpy(), ref["vdes"], rtol=1e-7, atol=1e-7, err_msg=f"vdes tick {k}")
        np.testing.assert_allclose(res["tau_ff"].cpu().numpy(), ref["tau_ff"], rtol=1e-5, atol=2e-4, err_msg=f"tau tick {k}")


def test_graph_replay_of_one_mpc_period_equals_tick_by_tick():
    import torch

    from qrw_b200 import workloads as wl
    from qrw_b200.controller import BatchedController

    B = 300
    seq = []
    W = wl.PlannerWorkload(B, 22, change_every=1000, yaw0=0.0, speed=0.5)
    for k in range(30):
        seq.append([torch.from_numpy(a).cuda() for a in W.tick(k)])
    A, G = BatchedController(B, h_ref=H_REF), BatchedController(B, h_ref=H_REF)
    for k in range(10):
        A.tick(k, *seq[k])
        G.tick(k, *seq[k])
    G.capture_period()
    taus = []
    for period in (1, 2):
        for i in range(10):
            k = 10 * period + i
            A.tick(k, *seq[k])
            taus.append(A.results()["tau_ff"].clone())
            G.in_q7[i].copy_(seq[k][0]), G.in_hv[i].copy_(seq[k][1]), G.in_vref[i].copy_(seq[k][2]), G.in_code[i].copy_(seq[k][3])
        G.replay_period()
        torch.cuda.synchronize()
        for i in range(10):
            assert torch.equal(G.out_tau[i], taus[10 * (period - 1) + i]), (period, i)
    assert torch.equal(G.mpc_result, A.mpc_result)
