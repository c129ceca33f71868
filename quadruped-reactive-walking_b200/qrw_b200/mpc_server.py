"""One GPU serving the MPC of many robot processes: the asynchronous mode of the reference's
``MPC_Wrapper`` (/root/reference/scripts/MPC_Wrapper.py:47-52, 150-264) with its child process replaced by a shared server.

* ``SharedBlock`` -- one named shared-memory block with R slots; slot r holds exactly the reference's shared objects:
  ``newData`` / ``newResult`` ('b' Values -> int8), ``dataIn`` (1 + 12 (n+1) + 12 N_gait doubles: ``k / k_mpc | xref C order |
  fsteps C order``, compress_dataIn :227-243), ``dataOut`` (24 n doubles, F order, convert_dataOut :260-264), plus one
  ``running`` flag for the block (:52, :296-306).
* ``MPCClient`` -- what a robot's control process holds instead of ``MPC_Wrapper(..., multiprocessing=True)``: same
  ``solve(k, xref, fsteps, gait)`` / ``get_latest_result()`` / ``stop_parallel_loop()`` and the same result buffering
  (default first result, roll / refill of ``last_available_result``, :83-126).  Pure numpy: no CUDA in the robot process.
* ``MPCServer`` -- owns the ``BatchedMPC`` of R robots and runs the child's loop (``create_MPC_asynchronous``, :170-225) for
  all slots natively (``qrw_mpc_server_loop_host``): slots with new data are claimed and solved in ONE launch of the
  per-robot-indexed MPC kernel; robots that posted nothing keep their solver state untouched.
"""
import ctypes as C
from multiprocessing import shared_memory

import numpy as np

from .mpc_wrapper import quaternion_to_rpy


def slot_sizes(n_steps=16, N_gait=20):
    return 1 + 12 * (n_steps + 1) + 12 * N_gait, 24 * n_steps


class SharedBlock:
    """Layout (bytes): [new_data R int8 | new_result R int8 | running 1 int8 | pad to 64] [data_in R x per_in f64] [data_out R x per_out f64]"""

    def __init__(self, name=None, robots=1, n_steps=16, N_gait=20, create=False):
        self.R, self.n_steps, self.N_gait = int(robots), int(n_steps), int(N_gait)
        self.per_in, self.per_out = slot_sizes(n_steps, N_gait)
        self.hdr = (2 * self.R + 1 + 63) // 64 * 64
        size = self.hdr + 8 * self.R * (self.per_in + self.per_out)
        self.shm = shared_memory.SharedMemory(name=name, create=create, size=size if create else 0)
        if not create and self.shm.size < size:
            raise ValueError("shared block is smaller than the requested layout")
        buf = self.shm.buf
        self.new_data = np.ndarray((self.R,), np.int8, buf, 0)
        self.new_result = np.ndarray((self.R,), np.int8, buf, self.R)
        self.running = np.ndarray((1,), np.int8, buf, 2 * self.R)
        self.data_in = np.ndarray((self.R, self.per_in), np.float64, buf, self.hdr)
        self.data_out = np.ndarray((self.R, self.per_out), np.float64, buf, self.hdr + 8 * self.R * self.per_in)
        if create:
            self.shm.buf[:size] = bytes(size)
            self.running[0] = 1
        self.name = self.shm.name

    def close(self, unlink=False):
        self.new_data = self.new_result = self.running = self.data_in = self.data_out = None
        self.shm.close()
        if unlink:
            self.shm.unlink()


def compress_data_in(k, k_mpc, xref, fsteps):
    """MPC_Wrapper.compress_dataIn (:227-243): [k / k_mpc | xref.ravel() | fsteps.ravel()] with NaN footsteps zeroed."""
    fsteps = np.array(fsteps, dtype=np.float64)
    fsteps[np.isnan(fsteps)] = 0.0
    return np.concatenate([[k / k_mpc], np.asarray(xref, np.float64).ravel(), fsteps.ravel()], axis=0)


class MPCClient:
    """Robot-side end of one slot; mirrors MPC_Wrapper in multiprocessing mode (the server plays the child process)."""

    def __init__(self, block, slot, k_mpc, q_init):
        self.block = block if isinstance(block, SharedBlock) else SharedBlock(**block)
        self.slot, self.k_mpc, self.n_steps = int(slot), k_mpc, self.block.n_steps
        self.not_first_iter = False
        q_init = np.asarray(q_init, dtype=np.float64).reshape(-1)
        x_init = np.zeros(12)
        x_init[0:3] = q_init[0:3]
        x_init[3:6] = quaternion_to_rpy(q_init[3:7])
        self.last_available_result = np.zeros((24, self.n_steps))
        self.last_available_result[:, 0] = np.hstack((x_init, np.array([0.0, 0.0, 8.0] * 4)))  # :64-71

    def solve(self, k, xref, fsteps, gait):
        b, s = self.block, self.slot
        b.data_in[s, :] = compress_data_in(k, self.k_mpc, xref, fsteps)  # run_MPC_asynchronous :166-167
        b.new_data[s] = 1
        n = self.n_steps
        if k > 2:  # :89-90
            self.last_available_result[12:(12 + n), :] = np.roll(self.last_available_result[12:(12 + n), :], -1, axis=1)
        pt = 0
        while np.any(gait[pt, :]):
            pt += 1
        if k > 2 and not np.array_equal(gait[0, :], gait[pt - 1, :]):  # :92-102
            F = 9.81 * 2.5 / np.sum(gait[pt - 1, :])
            self.last_available_result[12:, n - 1] = np.zeros(12)
            for i in range(4):
                if gait[pt - 1, i] == 1:
                    self.last_available_result[12 + 3 * i + 2, n - 1] = F
        return 0

    def get_latest_result(self):
        if self.not_first_iter:  # :110-119
            b, s = self.block, self.slot
            if b.new_result[s]:
                b.new_result[s] = 0
                self.last_available_result = np.array(b.data_out[s]).reshape((24, -1), order="F")  # convert_dataOut
            return self.last_available_result
        self.not_first_iter = True
        return self.last_available_result

    def stop_parallel_loop(self):  # :296-306 (stops the shared server: only the last robot standing should call it)
        self.block.running[0] = 0
        return 0


class MPCServer:
    def __init__(self, block, dt=0.02, T_gait=0.32, device=0):
        from ._lib import check, load_library
        from .batched import BatchedMPC

        self._L, self._check = load_library(), check
        self.block = block
        self.mpc = BatchedMPC(block.R, dt, block.n_steps, T_gait, block.N_gait, device)
        h = C.c_void_p()
        check(self._L.qrw_mpc_server_create(C.byref(h), self.mpc._h))
        self._h = h
        pi, po = C.c_int(), C.c_int()
        check(self._L.qrw_mpc_server_slot_doubles(h, C.byref(pi), C.byref(po)))
        assert (pi.value, po.value) == (block.per_in, block.per_out), "wire format mismatch between block and library"

    def _ptrs(self):
        b = self.block
        return b.data_in.ctypes.data, b.new_data.ctypes.data, b.data_out.ctypes.data, b.new_result.ctypes.data

    def poll(self):
        """One pass over the slots; returns how many robots were served."""
        n = C.c_int()
        self._check(self._L.qrw_mpc_server_poll_host(self._h, *self._ptrs(), C.byref(n)))
        return n.value

    def serve(self, idle_us=50, max_rounds=0):
        """`while running.value` (MPC_Wrapper.py:186); returns the number of solves done."""
        tot = C.c_longlong()
        self._check(self._L.qrw_mpc_server_loop_host(self._h, *self._ptrs(), self.block.running.ctypes.data, int(idle_us), int(max_rounds),
                                                     C.byref(tot)))
        return tot.value

    def close(self):
        if getattr(self, "_h", None):
            self._L.qrw_mpc_server_destroy(self._h)
            self._h = None
            self.mpc.close()
