"""Batched input producers on the GPU: Gait, FootstepPlanner, FootTrajectoryGenerator, StatePlanner and the feet
commands of the control loop, for B robots per object (C ABI ``qrw_planner_*``, csrc/planner_core.cuh).

Reference: /root/reference/src/Gait.cpp, FootstepPlanner.cpp, FootTrajectoryGenerator.cpp, StatePlanner.cpp and the
call order of /root/reference/scripts/Controller.py:222-240, 286-296.  numpy arrays go through the ``_host`` entry
points, torch CUDA tensors through the ``_device`` ones (asynchronous on torch's current stream).
"""
import ctypes as C

import numpy as np

from ._lib import check, load_library
from .batched import _DevView, _dev_ptr, _host, _is_cuda_tensor, _torch_stream

GAIT, FOOTSTEPS, FOOTTRAJ, STATE, FEETCMD, ALL = 1, 2, 4, 8, 16, 31

_OUT_NAMES = ("xref", "fsteps", "contacts", "o_target", "fpos", "fvel", "facc", "p_cmd", "v_cmd", "a_cmd")


class BatchedPlanner:
    """B robots' planners.  ``run(k, q7, h_v, v_ref, code)`` is one control tick k of Controller.py:222-236 (+ :286-296):
    q7 [B,7] = q.head(7) (x y z, quaternion x y z w), h_v [B,6], v_ref [B,6], code [B] joystick codes (None = 0)."""

    def __init__(self, batch, dt_mpc=0.02, dt_wbc=0.002, T_gait=0.32, T_mpc=0.32, N_gait=20, k_mpc=10, h_ref=0.2447495, device=0):
        self._L = load_library()
        self.B, self.device, self.N_gait = int(batch), int(device), int(N_gait)
        self.n_steps = int(round(T_mpc / dt_mpc))
        h = C.c_void_p()
        check(self._L.qrw_planner_create(C.byref(h), self.B, dt_mpc, dt_wbc, T_gait, T_mpc, self.N_gait, int(k_mpc), h_ref, self.device))
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            self._L.qrw_planner_destroy(self._h)
            self._h = None

    __del__ = close

    def configure(self, shoulders=None, ftg_target=None, ftg_position=None, max_height=0.05, lock_time=0.07):
        """The remaining initialize() arguments (3x4 matrices shared by the batch); resets every robot."""
        a = [None if m is None else np.ascontiguousarray(np.asarray(m, np.float64).reshape(3, 4).ravel(order="F"))
             for m in (shoulders, ftg_target, ftg_position)]
        check(self._L.qrw_planner_configure(self._h, *[None if x is None else x.ctypes.data for x in a], max_height, lock_time))

    def run(self, k, q7, h_v, v_ref, code=None, stages=ALL):
        B = self.B
        if _is_cuda_tensor(q7):
            cp = None
            if code is not None:
                assert code.is_cuda and code.numel() == B and code.dtype.itemsize == 4
                cp = C.c_void_p(code.data_ptr())
            check(self._L.qrw_planner_run_device(self._h, int(k), int(stages), _dev_ptr(q7, 7 * B), _dev_ptr(h_v, 6 * B),
                                                 _dev_ptr(v_ref, 6 * B), cp, _torch_stream()))
            return 0
        q, v, r = _host(q7, 7 * B), _host(h_v, 6 * B), _host(v_ref, 6 * B)
        cd = None if code is None else np.ascontiguousarray(np.broadcast_to(np.asarray(code, np.int32), (B,)))
        check(self._L.qrw_planner_run_host(self._h, int(k), int(stages), q.ctypes.data, v.ctypes.data, r.ctypes.data,
                                           None if cd is None else cd.ctypes.data))
        return 0

    def outputs(self, names=None):
        """Host copies: xref [B,12(n+1)], fsteps [B,12 N_gait], gait [B,4 N_gait], o_target / fpos / fvel / facc /
        p_cmd / v_cmd / a_cmd [B,12]."""
        B, N = self.B, self.N_gait
        shapes = dict(xref=12 * (self.n_steps + 1), fsteps=12 * N, gait=4 * N, o_target=12, fpos=12, fvel=12, facc=12, p_cmd=12,
                      v_cmd=12, a_cmd=12)
        order = ("xref", "fsteps", "gait", "o_target", "fpos", "fvel", "facc", "p_cmd", "v_cmd", "a_cmd")
        want = order if names is None else names
        out = {n: np.empty((B, shapes[n])) for n in want}
        check(self._L.qrw_planner_outputs_host(self._h, *[out[n].ctypes.data if n in out else None for n in order]))
        return out

    def output_tensors(self):
        """Zero-copy torch views of the device outputs (dict keyed like ``outputs`` with `contacts` [B,4])."""
        import torch

        ps = [C.c_void_p() for _ in _OUT_NAMES]
        check(self._L.qrw_planner_outputs_device(self._h, *[C.byref(p) for p in ps]))
        widths = dict(xref=12 * (self.n_steps + 1), fsteps=12 * self.N_gait, contacts=4)
        return {n: torch.as_tensor(_DevView(p.value, (self.B, widths.get(n, 12))), device=f"cuda:{self.device}")
                for n, p in zip(_OUT_NAMES, ps)}

    def state(self):
        B, N = self.B, self.N_gait
        s = dict(past=np.empty((B, 4 * N)), desired=np.empty((B, 4 * N)), cur_fs=np.empty((B, 12)), t0s=np.empty((B, 4)),
                 t_swing=np.empty((B, 4)))
        fl = np.empty((B, 2), np.int32)
        check(self._L.qrw_planner_state_host(self._h, *[s[k].ctypes.data for k in ("past", "desired", "cur_fs", "t0s", "t_swing")],
                                             fl.ctypes.data))
        s["flags"] = fl
        return s

    def set_target(self, o_target):
        t = _host(o_target, 12 * self.B)
        check(self._L.qrw_planner_set_target_host(self._h, t.ctypes.data))
