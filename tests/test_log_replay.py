"""Log replay: a control log in the reference's own format (LoggerControl.saveAll,
scripts/LoggerControl.py:551-628) is written by the ORACLE chain (planners -> MPC -> whole-body tick composed as
scripts/Controller.py:222-310 does) and replayed through the CUDA path.  CPU: format / history handling.  GPU: parity."""
import numpy as np
import pytest

from oracle import bindings as ob
from qrw_b200 import workloads as wl

H_REF = 0.2447495


def oracle_log(seed, T, k_mpc=10):
    """One robot's log with the LoggerControl keys the replay reads, produced by the oracle."""
    W = wl.PlannerWorkload(1, seed, change_every=10 ** 6, yaw0=0.0, speed=0.5)
    P, M, Q = ob.OraclePlanner(1, h_ref=H_REF), ob.OracleMPC(1), ob.OracleWBC(1)
    n, N = 16, 20
    L = dict(joy_v_ref=np.zeros((T, 6)), loop_o_q_int=np.zeros((T, 19)), loop_h_v=np.zeros((T, 18)), planner_xref=np.zeros((T, 12, n + 1)),
             planner_fsteps=np.zeros((T, N, 12)), planner_gait=np.zeros((T, N, 4)), planner_goals=np.zeros((T, 3, 4)),
             planner_vgoals=np.zeros((T, 3, 4)), planner_agoals=np.zeros((T, 3, 4)), planner_h_ref=np.full(T, H_REF),
             mpc_x_f=np.zeros((T, 24, n)), wbc_x_f=np.zeros((T, 24)), wbc_q_des=np.zeros((T, 12)), wbc_v_des=np.zeros((T, 12)),
             wbc_tau_ff=np.zeros((T, 12)), wbc_f_ctc=np.zeros((T, 12)), wbc_feet_pos_target=np.zeros((T, 3, 4)),
             wbc_feet_vel_target=np.zeros((T, 3, 4)), wbc_feet_acc_target=np.zeros((T, 3, 4)))
    x_f = np.zeros((24, n))
    x_f[:, 0] = np.hstack(([0, 0, H_REF, 0, 0, 0], np.zeros(6), [0.0, 0.0, 8.0] * 4))  # MPC_Wrapper.py:64-71
    first, f_applied = True, None
    qdes, vdes, pp, vp = wl.Q_INIT[None, :].copy(), np.zeros((1, 12)), np.zeros((1, 12)), np.zeros((1, 12))
    for k in range(T):
        q7, hv, vr, code = W.tick(k)
        P.run(k, q7, hv, vr, code, stages=15)
        o = P.outputs()
        if k % k_mpc == 0:
            M.run(k, o["xref"], o["fsteps"], 1)
            f_applied = M.get_latest_result().reshape(n, 24).T.copy()
        x_f_mpc = x_f if first else f_applied  # MPC_Wrapper.get_latest_result
        first = False
        xref = o["xref"].reshape(n + 1, 12).T
        x_f_wbc = x_f_mpc[:, 0].copy()  # Controller.py:258-269
        x_f_wbc[0:6] = [0.002 * xref[6, 1], 0.002 * xref[7, 1], H_REF, 0.0, 0.0, 0.002 * xref[11, 1]]
        x_f_wbc[6:12] = xref[6:, 1]
        p, v, a = ob.feet_commands(o["fpos"], o["fvel"], o["facc"], q7, vr, pp, vp, H_REF)
        pp, vp = p, v
        q19, dq18 = np.zeros((1, 19)), np.zeros((1, 18))
        q19[:, 2], q19[:, 6], q19[:, 7:] = H_REF, 1.0, qdes
        dq18[:, :6], dq18[:, 6:] = vr, vdes
        gait = o["gait"].reshape(4, N).T
        ref = ob.wbc_tick(Q, q19, dq18, x_f_wbc[None, 12:], gait[None, 0, :].copy(), p, v, a)
        qdes, vdes = ref["qdes"], ref["vdes"]
        L["joy_v_ref"][k], L["loop_o_q_int"][k, :7], L["loop_h_v"][k, :6] = vr[0], q7[0], hv[0]
        L["planner_xref"][k], L["planner_fsteps"][k], L["planner_gait"][k] = xref, o["fsteps"].reshape(12, N).T, gait
        for key, src in (("planner_goals", o["fpos"]), ("planner_vgoals", o["fvel"]), ("planner_agoals", o["facc"]),
                         ("wbc_feet_pos_target", p), ("wbc_feet_vel_target", v), ("wbc_feet_acc_target", a)):
            L[key][k] = src.reshape(4, 3).T
        L["mpc_x_f"][k], L["wbc_x_f"][k] = x_f_mpc, x_f_wbc
        L["wbc_q_des"][k], L["wbc_v_des"][k], L["wbc_tau_ff"][k], L["wbc_f_ctc"][k] = qdes[0], vdes[0], 0.8 * ref["tau_ff"][0], ref["f_res"][0]
    return L


def test_log_format_and_missing_keys(tmp_path):
    from qrw_b200.log_replay import LogReplay, _cm

    L = oracle_log(3, 12)
    path = tmp_path / "data_2021_01_01_00_00.npz"
    np.savez(path, **L)  # what LoggerControl.saveAll writes (np.savez, same keys)
    R = LogReplay([str(path), L])
    assert (R.B, R.T, R.n_steps, R.N_gait) == (2, 12, 16, 20)
    m = np.arange(24.0).reshape(1, 4, 6)
    assert (_cm(m)[0] == m[0].ravel(order="F")).all()
    bad = dict(L)
    del bad["mpc_x_f"]
    with pytest.raises(KeyError, match="mpc_x_f"):
        LogReplay([bad])._need(("planner_xref", "mpc_x_f"))
    # the log is self-consistent the way the reference's is: tick 0 carries the wrapper's default block
    assert (L["mpc_x_f"][0][12:, 0] == np.array([0.0, 0.0, 8.0] * 4)).all() and (L["mpc_x_f"][1] == L["mpc_x_f"][9]).all()
    assert (L["wbc_x_f"][5][12:] == L["mpc_x_f"][5][12:, 0]).all()


@pytest.mark.gpu
def test_replay_of_oracle_logs_on_the_cuda_path():
    from qrw_b200.log_replay import LogReplay

    R = LogReplay([oracle_log(s, 45) for s in (1, 2, 3)])
    m = R.replay_mpc()
    assert m["ticks"] == [0, 10, 20, 30, 40]
    assert max(m["max_abs_force"]) < 1e-4 and max(m["max_abs_state"]) < 1e-4
    p = R.replay_planners()
    assert p["gait"] == 0.0 and max(p[k] for k in ("xref", "fsteps", "goals", "vgoals", "agoals")) < 1e-8
    w = R.replay_wbc()
    assert w["f_ctc"] < 1e-4 and w["q_des"] < 1e-8 and w["v_des"] < 1e-7 and w["tau_ff"] < 2e-4
