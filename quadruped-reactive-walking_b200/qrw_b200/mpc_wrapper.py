"""Host-side mirror of the reference's ``MPC_Wrapper`` (scripts/MPC_Wrapper.py:21-148), synchronous mode, for one
robot (same constructor and method names, so scripts/Controller.py:144,246-253 can use it unchanged) and for a batch.

What the wrapper adds around ``MPC.run`` / ``get_latest_result`` (all host logic, no arithmetic of the solve):
  * the default result of the very first call: x_init | [0, 0, 8] x 4 in column 0 (MPC_Wrapper.py:64-71);
  * after every solve with k > 2: rows 12..23 of ``last_available_result`` rolled one column to the left
    (:89-90, written there as rows 12:(12+n_steps)) and, when the gait's last active row differs from its first,
    the last column refilled with the weight 9.81 * 2.5 shared by that row's stance feet (:92-102);
  * ``get_latest_result`` returns the default once, then the solver's result (:104-126).
The multiprocessing mode of the reference forks a child that builds its own MPC (:150-225); with this library
the child would create its CUDA context after the fork, which works, but the shared-memory transport is a
"next" item and is not mirrored here.
"""
import numpy as np


def quaternion_to_rpy(q):
    """scripts/utils_mpc.py quaternionToRPY: (x, y, z, w) -> roll, pitch, yaw."""
    x, y, z, w = [float(v) for v in q]
    roll = np.arctan2(2.0 * (w * x + y * z), 1.0 - 2.0 * (x * x + y * y))
    pitch = np.arcsin(np.clip(2.0 * (w * y - z * x), -1.0, 1.0))
    yaw = np.arctan2(2.0 * (w * z + x * y), 1.0 - 2.0 * (y * y + z * z))
    return np.array([roll, pitch, yaw])


class MPC_Wrapper:
    def __init__(self, mpc_type, dt, n_steps, k_mpc, T_gait, N_gait, q_init, multiprocessing=False, mpc=None):
        if multiprocessing:
            raise NotImplementedError("qrw_b200.MPC_Wrapper mirrors the synchronous mode only (SURVEY.md 8f-3)")
        if not mpc_type:
            raise NotImplementedError("the Crocoddyl back end (type_MPC = False) is out of scope")
        self.f_applied = np.zeros((12,))
        self.not_first_iter = False
        self.k_mpc, self.dt, self.n_steps, self.T_gait, self.N_gait = k_mpc, dt, int(n_steps), T_gait, int(N_gait)
        self.gait_memory = np.zeros(4)
        self.mpc_type, self.multiprocessing = mpc_type, False
        if mpc is None:
            import libquadruped_reactive_walking as MPC

            mpc = MPC.MPC(dt, self.n_steps, T_gait, self.N_gait)
        self.mpc = mpc
        q_init = np.asarray(q_init, dtype=np.float64).reshape(-1)
        x_init = np.zeros(12)
        x_init[0:3] = q_init[0:3]
        x_init[3:6] = quaternion_to_rpy(q_init[3:7])
        self.last_available_result = np.zeros((24, self.n_steps))
        self.last_available_result[:, 0] = np.hstack((x_init, np.array([0.0, 0.0, 8.0] * 4)))

    def solve(self, k, xref, fsteps, gait):
        self.run_MPC_synchronous(k, xref, fsteps)
        n = self.n_steps
        if k > 2:
            self.last_available_result[12:(12 + n), :] = np.roll(self.last_available_result[12:(12 + n), :], -1, axis=1)
        pt = 0
        while np.any(gait[pt, :]):
            pt += 1
        if k > 2 and not np.array_equal(gait[0, :], gait[pt - 1, :]):
            mass = 2.5
            nb_ctc = np.sum(gait[pt - 1, :])
            F = 9.81 * mass / nb_ctc
            self.last_available_result[12:, n - 1] = np.zeros(12)
            for i in range(4):
                if gait[pt - 1, i] == 1:
                    self.last_available_result[12 + 3 * i + 2, n - 1] = F
        return 0

    def get_latest_result(self):
        if self.not_first_iter:
            return self.f_applied
        self.not_first_iter = True
        return self.last_available_result

    def run_MPC_synchronous(self, k, xref, fsteps):
        self.mpc.run(int(k), xref.copy(), fsteps.copy())
        self.f_applied = self.mpc.get_latest_result()

    def stop_parallel_loop(self):  # MPC_Wrapper.py:296-306, nothing to stop in synchronous mode
        return 0


class BatchedMPCWrapper:
    """The same semantics for B robots on one ``BatchedMPC``: solve(k, xref [B,204], fsteps [B,240]) and
    get_latest_result() -> [B, 24, n_steps] (the default [x_init | 0 0 8 ...] block on the first call)."""

    def __init__(self, batch, dt, n_steps, k_mpc, T_gait, N_gait, q_init, mpc=None, device=0):
        from .batched import BatchedMPC, unpack_cols

        self._unpack = unpack_cols
        self.B, self.n_steps, self.k_mpc = int(batch), int(n_steps), k_mpc
        self.mpc = mpc if mpc is not None else BatchedMPC(batch, dt, n_steps, T_gait, N_gait, device)
        q_init = np.asarray(q_init, dtype=np.float64).reshape(-1, 19) if np.size(q_init) >= 19 else np.asarray(q_init).reshape(-1, 7)
        q_init = np.broadcast_to(q_init, (self.B, q_init.shape[1]))
        self.last_available_result = np.zeros((self.B, 24, self.n_steps))
        for b in range(self.B):
            self.last_available_result[b, 0:3, 0] = q_init[b, 0:3]
            self.last_available_result[b, 3:6, 0] = quaternion_to_rpy(q_init[b, 3:7])
        self.last_available_result[:, 12:, 0] = np.array([0.0, 0.0, 8.0] * 4)
        self.not_first_iter = False
        self.f_applied = None

    def solve(self, k, xref, fsteps):
        self.mpc.run(int(k), xref, fsteps)
        self.f_applied = self._unpack(self.mpc.get_latest_result(), 24, self.n_steps)
        return 0

    def get_latest_result(self):
        if self.not_first_iter:
            return self.f_applied
        self.not_first_iter = True
        return self.last_available_result
