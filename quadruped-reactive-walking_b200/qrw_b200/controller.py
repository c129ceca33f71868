"""The control tick of /root/reference/scripts/Controller.py:222-303 for B robots, entirely in HBM -- host-side view of the
C ABI ``qrw_tick_*`` (include/qrw_b200.h, csrc/tick_abi.cuh):

    planners (Gait, FootstepPlanner, FootTrajectoryGenerator, StatePlanner, feet commands)
    -> every k_mpc-th tick: MPC::run on the planners' xref / fsteps
    -> whole-body tick with f_cmd = x_f_mpc[12:24, 0], contacts = gait row 0, feet commands

Nothing crosses PCIe inside a tick except what the caller feeds (base pose q.head(7), base velocity h_v, velocity reference,
joystick code: 156 B per robot) and reads (tau_ff, q_des, v_des: 288 B per robot).  The composition, the default forces of
the very first tick (MPC_Wrapper.py:64-71, :122-126) and the CUDA graph of one MPC period live in the library; this class
only converts between torch / numpy arrays and pointers, so a C++ caller reaches exactly the same path.
"""
import ctypes as C

import numpy as np

from ._lib import check, load_library
from .batched import BatchedMPC, BatchedWbcTick, _DevView, _torch_stream
from .planner import BatchedPlanner
from .workloads import Q_INIT


def _borrow(cls, handle, **attrs):
    """A Batched* view of a sub-handle owned by the tick handle (never destroyed through the view)."""
    view = type("Borrowed" + cls.__name__, (cls,), {"close": lambda self: None, "__del__": lambda self: None})
    o = view.__new__(view)
    o._L, o._h = load_library(), handle
    o.__dict__.update(attrs)
    return o


class BatchedController:
    def __init__(self, batch, dt_wbc=0.002, dt_mpc=0.02, T_gait=0.32, T_mpc=0.32, N_gait=20, h_ref=0.2447495, q_init=Q_INIT, device=0):
        import torch

        self._L = load_library()
        self.B, self.device = int(batch), int(device)
        self.k_mpc = int(round(dt_mpc / dt_wbc))
        self.n_steps = int(round(T_mpc / dt_mpc))
        self.N_gait = int(N_gait)
        self.h_ref = float(h_ref)
        qi = np.ascontiguousarray(np.asarray(q_init, np.float64).reshape(12))
        h = C.c_void_p()
        check(self._L.qrw_tick_create(C.byref(h), self.B, float(dt_wbc), float(dt_mpc), float(T_gait), float(T_mpc), self.N_gait,
                                      self.h_ref, qi.ctypes.data, self.device))
        self._h = h
        self.dev = torch.device("cuda", self.device)
        pl, mp, wb = C.c_void_p(), C.c_void_p(), C.c_void_p()
        check(self._L.qrw_tick_parts(self._h, C.byref(pl), C.byref(mp), C.byref(wb)))
        # views of the parts (getters, info, state snapshots); the tick handle owns them
        self.planner = _borrow(BatchedPlanner, pl, B=self.B, device=self.device, N_gait=self.N_gait, n_steps=self.n_steps)
        self.mpc = _borrow(BatchedMPC, mp, B=self.B, n=self.n_steps, N_gait=self.N_gait, device=self.device)
        self.wbc = _borrow(BatchedWbcTick, wb, B=self.B, device=self.device)
        self.mpc_result = self.mpc.result_tensor()  # [B, 24 * n_steps]
        self._captured = False

    def close(self):
        if getattr(self, "_h", None):
            for o in (self.planner, self.mpc, self.wbc):
                o._h = None
            self._L.qrw_tick_destroy(self._h)
            self._h = None

    __del__ = close

    def tick(self, k, q7, h_v, v_ref, code=None):
        """One control tick.  CUDA tensors ([B,7], [B,6], [B,6] float64, [B] int32 or None) run asynchronously on torch's
        current stream; numpy arrays go through the host entry point (H2D, tick, synchronised)."""
        if hasattr(q7, "is_cuda"):
            cp = None if code is None else C.c_void_p(code.data_ptr())
            check(self._L.qrw_tick_run_device(self._h, int(k), C.c_void_p(q7.data_ptr()), C.c_void_p(h_v.data_ptr()),
                                              C.c_void_p(v_ref.data_ptr()), cp, _torch_stream()))
        else:
            a = [np.ascontiguousarray(x, np.float64) for x in (q7, h_v, v_ref)]
            cd = None if code is None else np.ascontiguousarray(np.broadcast_to(np.asarray(code, np.int32), (self.B,)))
            check(self._L.qrw_tick_run_host(self._h, int(k), *[x.ctypes.data for x in a], None if cd is None else cd.ctypes.data))
        return 0

    def results(self):
        """Zero-copy torch views: tau_ff, q_des, v_des, f_res ([B,12] each; Controller.py:306-310 scales tau_ff by 0.8)."""
        import torch

        ps = [C.c_void_p() for _ in range(4)]
        check(self._L.qrw_tick_result_device(self._h, *[C.byref(p) for p in ps]))
        return {n: torch.as_tensor(_DevView(p.value, (self.B, 12)), device=self.dev) for n, p in zip(("tau_ff", "q_des", "v_des", "f_res"), ps)}

    def results_host(self):
        outs = [np.empty((self.B, 12)) for _ in range(4)]
        check(self._L.qrw_tick_result_host(self._h, *[o.ctypes.data for o in outs]))
        return dict(zip(("tau_ff", "q_des", "v_des", "f_res"), outs))

    # ---- one MPC period as a CUDA graph (captured inside the library)
    def capture_period(self):
        """Records ticks k_mpc .. 2 k_mpc - 1 into a CUDA graph; returns the staging tensors in_q7 / in_hv / in_vref
        [k_mpc, B, .] and in_code [k_mpc, B].  Must be called after at least one tick."""
        import torch

        check(self._L.qrw_tick_capture_period(self._h))
        ps = [C.c_void_p() for _ in range(7)]
        check(self._L.qrw_tick_period_buffers(self._h, *[C.byref(p) for p in ps]))
        K, B = self.k_mpc, self.B

        def view(p, w, typ="<f8"):
            v = _DevView(p.value, (K, B, w) if w else (K, B))
            v.__cuda_array_interface__["typestr"] = typ
            return torch.as_tensor(v, device=self.dev)

        self.in_q7, self.in_hv, self.in_vref = view(ps[0], 7), view(ps[1], 6), view(ps[2], 6)
        self.in_code = view(ps[3], 0, "<i4")
        self.out_tau, self.out_qdes, self.out_vdes = view(ps[4], 12), view(ps[5], 12), view(ps[6], 12)
        self._captured = True
        return self.in_q7, self.in_hv, self.in_vref, self.in_code

    def replay_period(self):
        check(self._L.qrw_tick_replay_period(self._h, _torch_stream()))

    def period_host(self, q7, h_v, v_ref, code=None, out=None):
        """One MPC period from host arrays [k_mpc, B, .]: H2D, one graph launch, D2H of tau_ff / q_des / v_des [k_mpc, B, 12]."""
        K, B = self.k_mpc, self.B
        a = [np.ascontiguousarray(x, np.float64) for x in (q7, h_v, v_ref)]
        cd = None if code is None else np.ascontiguousarray(code, np.int32)
        if out is None:
            out = [np.empty((K, B, 12)) for _ in range(3)]
        check(self._L.qrw_tick_period_host(self._h, *[x.ctypes.data for x in a], None if cd is None else cd.ctypes.data,
                                           *[o.ctypes.data for o in out]))
        return out

    # ---- state snapshot (benchmark replay / checkpoint): MPC solver state only; planners and qdes are cheap to re-run
    def snapshot(self):
        import torch

        buf = torch.empty(self.mpc.state_bytes(), dtype=torch.uint8, device=self.dev)
        self.mpc.state_save(buf)
        return buf

    def restore(self, buf):
        self.mpc.state_load(buf)
