"""Replay of the reference's control logs through the CUDA path.

``LoggerControl.saveAll`` (/root/reference/scripts/LoggerControl.py:551-628) writes one ``.npz`` per run with, per control
tick i: ``planner_xref`` [12, n+1], ``planner_fsteps`` [N_gait, 12], ``planner_gait`` [N_gait, 4], ``planner_goals`` /
``vgoals`` / ``agoals`` [3, 4] (world frame), ``mpc_x_f`` [24, n] (the MPC result in use at tick i), ``wbc_x_f`` [24],
``wbc_feet_pos_target`` / ``vel_target`` / ``acc_target`` [3, 4] (base frame: what ``wbc_controller.compute`` received),
``wbc_q_des`` / ``wbc_v_des`` / ``wbc_tau_ff`` / ``wbc_f_ctc`` [12], ``joy_v_ref`` [6], ``loop_o_q_int`` [19], ``loop_h_v`` [18].
Every array is exactly an input or an output of one of the replaced functions, so a log of the real reference is a parity
vector for this library: ``LogReplay`` feeds the logged inputs of B logs (one robot each, equal length) through
``BatchedMPC`` / ``BatchedPlanner`` / ``BatchedWbcTick`` with the logged call history and reports the distance to the
logged outputs.  No reference log ships with the reference repository; tests build one with the oracle.
"""
import numpy as np

from .workloads import Q_INIT

MPC_KEYS = ("planner_xref", "planner_fsteps", "mpc_x_f")
PLANNER_KEYS = ("loop_o_q_int", "loop_h_v", "joy_v_ref", "planner_xref", "planner_fsteps", "planner_gait", "planner_goals", "planner_vgoals",
                "planner_agoals")
WBC_KEYS = ("wbc_x_f", "planner_gait", "wbc_feet_pos_target", "wbc_feet_vel_target", "wbc_feet_acc_target", "joy_v_ref", "wbc_q_des",
            "wbc_v_des", "wbc_tau_ff", "wbc_f_ctc")


def _cm(a):
    """[B, r, c] matrices -> [B, r*c] column-major (the C ABI's per-instance layout)."""
    a = np.asarray(a, dtype=np.float64)
    return np.ascontiguousarray(a.transpose(0, 2, 1)).reshape(a.shape[0], -1)


class LogReplay:
    def __init__(self, logs, k_mpc=10, dt_wbc=0.002, dt_mpc=0.02, T_gait=0.32, device=0):
        """logs: list of ``.npz`` paths or of dicts with the LoggerControl keys (one robot per log)."""
        self.logs = [dict(np.load(l)) if isinstance(l, (str, bytes)) or hasattr(l, "__fspath__") else l for l in logs]
        if not self.logs:
            raise ValueError("no logs")
        self.B, self.k_mpc, self.dt_wbc, self.dt_mpc, self.T_gait, self.device = len(self.logs), int(k_mpc), dt_wbc, dt_mpc, T_gait, device
        self.T = min(len(l["planner_xref"]) for l in self.logs)
        x = self.logs[0]["planner_xref"]
        self.n_steps, self.N_gait = x.shape[2] - 1, self.logs[0]["planner_fsteps"].shape[1]

    def _need(self, keys):
        for i, l in enumerate(self.logs):
            miss = [k for k in keys if k not in l]
            if miss:
                raise KeyError(f"log {i} lacks {miss} (LoggerControl.saveAll keys)")

    def _stack(self, key, i):
        return np.stack([np.asarray(l[key][i], dtype=np.float64) for l in self.logs])

    # ------------------------------------------------------------------ MPC::run, every k_mpc-th tick
    def replay_mpc(self, ticks=None):
        """Re-solves every logged MPC call (ticks i with i % k_mpc == 0, history from i = 0) and compares the result with
        ``mpc_x_f`` of the first tick that used it (i itself in synchronous mode, except i = 0 where MPC_Wrapper returns its
        default block, MPC_Wrapper.py:64-71,122-126).  Returns dict(ticks, max_abs_force, max_abs_state, iters)."""
        from .batched import BatchedMPC

        self._need(MPC_KEYS)
        T = self.T if ticks is None else min(self.T, ticks)
        mpc = BatchedMPC(self.B, self.dt_mpc, self.n_steps, self.T_gait, self.N_gait, self.device)
        out = dict(ticks=[], max_abs_force=[], max_abs_state=[], iters=[])
        for i in range(0, T, self.k_mpc):
            mpc.run(i, _cm(self._stack("planner_xref", i)), _cm(self._stack("planner_fsteps", i)))
            res = mpc.get_latest_result().reshape(self.B, self.n_steps, 24).transpose(0, 2, 1)  # [B, 24, n]
            j = i if i > 0 else 1  # tick 0 logged the wrapper's default block; tick 1 is the first to see the solve of tick 0
            if j >= T:
                break
            want = self._stack("mpc_x_f", j)
            out["ticks"].append(i)
            out["max_abs_force"].append(float(np.abs(res[:, 12:] - want[:, 12:]).max()))
            out["max_abs_state"].append(float(np.abs(res[:, :12] - want[:, :12]).max()))
            out["iters"].append(mpc.info()["iters"].copy())
        mpc.close()
        return out

    # ------------------------------------------------------------------ Gait / FootstepPlanner / FootTrajectoryGenerator / StatePlanner
    def replay_planners(self, h_ref=None, ticks=None, codes=None):
        """Feeds q.head(7), h_v and v_ref of every tick to the planner kernel and compares xref, fsteps, gait and the foot
        trajectory with the log.  The joystick code is not logged: pass ``codes`` [T, B] if the run changed gait."""
        from .planner import BatchedPlanner

        self._need(PLANNER_KEYS)
        T = self.T if ticks is None else min(self.T, ticks)
        if h_ref is None:
            h_ref = float(self.logs[0]["planner_h_ref"][0]) if "planner_h_ref" in self.logs[0] else 0.2447495
        pl = BatchedPlanner(self.B, self.dt_mpc, self.dt_wbc, self.T_gait, self.dt_mpc * self.n_steps, self.N_gait, self.k_mpc, h_ref, self.device)
        worst = dict(xref=0.0, fsteps=0.0, gait=0.0, goals=0.0, vgoals=0.0, agoals=0.0)
        for i in range(T):
            q7 = self._stack("loop_o_q_int", i)[:, :7]
            hv = self._stack("loop_h_v", i)[:, :6]
            vr = self._stack("joy_v_ref", i)[:, :6]
            pl.run(i, q7, hv, vr, None if codes is None else codes[i], stages=15)
            o = pl.outputs(("xref", "fsteps", "gait", "fpos", "fvel", "facc"))
            cmp = dict(xref=(o["xref"], _cm(self._stack("planner_xref", i))), fsteps=(o["fsteps"], _cm(self._stack("planner_fsteps", i))),
                       gait=(o["gait"], _cm(self._stack("planner_gait", i))), goals=(o["fpos"], _cm(self._stack("planner_goals", i))),
                       vgoals=(o["fvel"], _cm(self._stack("planner_vgoals", i))), agoals=(o["facc"], _cm(self._stack("planner_agoals", i))))
            for k, (got, want) in cmp.items():
                worst[k] = max(worst[k], float(np.abs(got - want).max()))
        pl.close()
        return worst

    # ------------------------------------------------------------------ wbc_controller.compute, every tick
    def replay_wbc(self, h_ref=None, q_init=Q_INIT, ticks=None):
        """Re-runs the whole-body tick of every logged control tick with the logged f_cmd (``wbc_x_f[12:]``), contacts
        (``planner_gait[0]``) and feet commands, q_wbc / b_v rebuilt from the logged q_des / v_des of the previous tick
        (Controller.py:275-283), and compares q_des, v_des, f_ctc and tau_ff (the log holds 0.8 tau_ff, Controller.py:310)."""
        from .batched import BatchedWbcTick

        self._need(WBC_KEYS)
        T = self.T if ticks is None else min(self.T, ticks)
        if h_ref is None:
            h_ref = float(self.logs[0]["planner_h_ref"][0]) if "planner_h_ref" in self.logs[0] else 0.2447495
        B = self.B
        wbc = BatchedWbcTick(B, self.device)
        worst = dict(q_des=0.0, v_des=0.0, f_ctc=0.0, tau_ff=0.0)
        qd, vd = np.tile(np.asarray(q_init, np.float64), (B, 1)), np.zeros((B, 12))
        for i in range(T):
            q, dq = np.zeros((B, 19)), np.zeros((B, 18))
            q[:, 2], q[:, 6], q[:, 7:] = h_ref, 1.0, qd
            dq[:, :6], dq[:, 6:] = self._stack("joy_v_ref", i)[:, :6], vd
            wbc.compute(q, dq, self._stack("wbc_x_f", i)[:, 12:], self._stack("planner_gait", i)[:, 0, :], _cm(self._stack("wbc_feet_pos_target", i)),
                        _cm(self._stack("wbc_feet_vel_target", i)), _cm(self._stack("wbc_feet_acc_target", i)))
            want = dict(q_des=self._stack("wbc_q_des", i), v_des=self._stack("wbc_v_des", i), f_ctc=self._stack("wbc_f_ctc", i),
                        tau_ff=self._stack("wbc_tau_ff", i))
            got = dict(q_des=wbc.qdes, v_des=wbc.vdes, f_ctc=wbc.f_with_delta, tau_ff=0.8 * wbc.tau_ff)
            for k in worst:
                worst[k] = max(worst[k], float(np.abs(got[k] - want[k]).max()))
            qd, vd = want["q_des"], want["v_des"]  # the logged history, as the reference ran it
        wbc.close()
        return worst
