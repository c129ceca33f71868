"""``MPC_Wrapper`` semantics (reference scripts/MPC_Wrapper.py:21-148, synchronous mode) for a BATCH of robots.

For ONE robot nothing is mirrored here: the reference's own ``scripts/MPC_Wrapper.py`` runs unmodified on the drop-in module
``libquadruped_reactive_walking.py`` (its constructor, ``solve``, ``run_MPC_synchronous`` with ``self.mpc.run(np.int(k), ...)``,
``get_latest_result``), which ``tests/test_reference_scripts.py`` executes against the oracle.  ``BatchedMPCWrapper`` keeps the
two things the wrapper adds around ``MPC.run`` that matter to a caller of a batch -- the default result of the very first call
(``x_init | [0, 0, 8] x 4`` in column 0, MPC_Wrapper.py:64-71) and ``get_latest_result`` returning it once (:104-126).  The
asynchronous (multiprocessing) mode of the reference is served by ``qrw_b200/mpc_server.py`` (shared-memory wire format, one GPU
for many robot processes).
"""
import numpy as np


def quaternion_to_rpy(q):
    """scripts/utils_mpc.py quaternionToRPY: (x, y, z, w) -> roll, pitch, yaw."""
    x, y, z, w = [float(v) for v in q]
    roll = np.arctan2(2.0 * (w * x + y * z), 1.0 - 2.0 * (x * x + y * y))
    pitch = np.arcsin(np.clip(2.0 * (w * y - z * x), -1.0, 1.0))
    yaw = np.arctan2(2.0 * (w * z + x * y), 1.0 - 2.0 * (y * y + z * z))
    return np.array([roll, pitch, yaw])


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
