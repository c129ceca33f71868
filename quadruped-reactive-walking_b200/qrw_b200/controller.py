"""The control tick of /root/reference/scripts/Controller.py:222-303 for B robots, entirely in HBM:

    planners (Gait, FootstepPlanner, FootTrajectoryGenerator, StatePlanner, feet commands)   qrw_planner_run_device
    -> every k_mpc-th tick: MPC::run on the planners' xref / fsteps                            qrw_mpc_run_device
    -> whole-body tick with f_cmd = x_f_mpc[12:24, 0], contacts = gait row 0, feet commands    qrw_wbctick_run_ctrl_device

Nothing crosses PCIe inside a tick except what the caller feeds (base pose q.head(7), base velocity h_v, velocity reference,
joystick code: 156 B per robot) and reads (tau_ff, q_des, v_des: 288 B per robot).  The MPC wrapper semantics of the
synchronous mode are kept: the very first tick uses the default forces [0, 0, 8] per foot (MPC_Wrapper.py:64-71,
:122-126), every later tick the latest MPC result.  The static-gait branch of Controller.py:266-268 reads attributes that
do not exist in the reference (``self.planner``) and is not mirrored.

``capture_period()`` records the k_mpc ticks of one MPC period into a CUDA graph (the tick index enters the kernels only
through k % k_mpc and k != 0, so the graph of ticks k_mpc .. 2 k_mpc - 1 replays for every later period); the caller
writes the period's inputs into the staging tensors and launches the graph once per period.
"""
import ctypes as C

import numpy as np

from ._lib import check, load_library
from .batched import BatchedMPC, BatchedWbcTick, _torch_stream
from .planner import ALL, BatchedPlanner
from .workloads import Q_INIT


class BatchedController:
    def __init__(self, batch, dt_wbc=0.002, dt_mpc=0.02, T_gait=0.32, T_mpc=0.32, N_gait=20, h_ref=0.2447495, q_init=Q_INIT, device=0):
        import torch

        self._L = load_library()
        self.B, self.device = int(batch), int(device)
        self.k_mpc = int(round(dt_mpc / dt_wbc))
        self.n_steps = int(round(T_mpc / dt_mpc))
        self.h_ref = float(h_ref)
        self.planner = BatchedPlanner(batch, dt_mpc, dt_wbc, T_gait, T_mpc, N_gait, self.k_mpc, h_ref, device)
        self.mpc = BatchedMPC(batch, dt_mpc, self.n_steps, T_gait, N_gait, device)
        self.wbc = BatchedWbcTick(batch, device)
        qd = np.ascontiguousarray(np.broadcast_to(np.asarray(q_init, np.float64).reshape(-1, 12), (self.B, 12)))
        check(self._L.qrw_wbctick_set_desired_host(self.wbc._h, qd.ctypes.data, None))  # Controller.py:153
        self.dev = torch.device("cuda", self.device)
        self.out = self.planner.output_tensors()
        self.mpc_result = self.mpc.result_tensor()  # [B, 24 * n_steps]
        self.default_forces = torch.tensor([0.0, 0.0, 8.0] * 4, dtype=torch.float64, device=self.dev).repeat(self.B, 1).contiguous()
        self.have_result = False  # MPC_Wrapper.not_first_iter
        self._graph = None

    def close(self):
        for o in (self.planner, self.mpc, self.wbc):
            o.close()

    def tick(self, k, q7, h_v, v_ref, code=None):
        """One control tick; all arguments are CUDA tensors ([B,7], [B,6], [B,6] float64, [B] int32 or None)."""
        self.planner.run(k, q7, h_v, v_ref, code, stages=ALL)
        if k % self.k_mpc == 0:
            self.mpc.run(k, self.out["xref"], self.out["fsteps"])  # MPC_Wrapper.solve -> MPC::run(k, xref, fsteps)
        if self.have_result:
            f_ptr, stride = self.mpc_result.data_ptr() + 12 * 8, 24 * self.n_steps
        else:
            f_ptr, stride = self.default_forces.data_ptr(), 12
            self.have_result = True
        o = self.out
        check(self._L.qrw_wbctick_run_ctrl_device(
            self.wbc._h, self.h_ref, C.c_void_p(v_ref.data_ptr()), C.c_void_p(f_ptr), stride, C.c_void_p(o["contacts"].data_ptr()),
            C.c_void_p(o["p_cmd"].data_ptr()), C.c_void_p(o["v_cmd"].data_ptr()), C.c_void_p(o["a_cmd"].data_ptr()), _torch_stream()))
        return 0

    def results(self):
        """Zero-copy torch views: tau_ff, q_des, v_des, f_res ([B,12] each; Controller.py:306-310 scales tau_ff by 0.8)."""
        return dict(zip(("tau_ff", "q_des", "v_des", "f_res"), self.wbc.result_tensors()))

    # ---- one MPC period as a CUDA graph
    def capture_period(self):
        """Staging tensors in_q7 / in_hv / in_vref [k_mpc, B, .] and in_code [k_mpc, B]; returns them.  Must be called
        after at least k_mpc ticks (the captured ticks are warm: k != 0)."""
        import torch

        assert self.have_result, "run the first period tick by tick before capturing"
        K, B = self.k_mpc, self.B
        self.in_q7 = torch.zeros(K, B, 7, dtype=torch.float64, device=self.dev)
        self.in_q7[:, :, 6] = 1.0
        self.in_hv = torch.zeros(K, B, 6, dtype=torch.float64, device=self.dev)
        self.in_vref = torch.zeros(K, B, 6, dtype=torch.float64, device=self.dev)
        self.in_code = torch.zeros(K, B, dtype=torch.int32, device=self.dev)
        # per-tick outputs of the period (the graph's results): [k_mpc, B, 12]
        self.out_tau = torch.zeros(K, B, 12, dtype=torch.float64, device=self.dev)
        self.out_qdes = torch.zeros_like(self.out_tau)
        self.out_vdes = torch.zeros_like(self.out_tau)
        res = self.results()
        snap = self.snapshot()
        s = torch.cuda.Stream(device=self.dev)
        s.wait_stream(torch.cuda.current_stream(self.dev))
        g = torch.cuda.CUDAGraph()
        with torch.cuda.stream(s):
            with torch.cuda.graph(g, stream=s):
                for i in range(K):
                    self.tick(K + i, self.in_q7[i], self.in_hv[i], self.in_vref[i], self.in_code[i])
                    self.out_tau[i].copy_(res["tau_ff"])
                    self.out_qdes[i].copy_(res["q_des"])
                    self.out_vdes[i].copy_(res["v_des"])
        torch.cuda.current_stream(self.dev).wait_stream(s)
        self._graph = g
        self.restore(snap)  # capture does not execute, but keep the contract explicit: state is what it was
        return self.in_q7, self.in_hv, self.in_vref, self.in_code

    def replay_period(self):
        self._graph.replay()

    # ---- state snapshot (benchmark replay / checkpoint): MPC solver state only; planners and qdes are cheap to re-run
    def snapshot(self):
        import torch

        buf = torch.empty(self.mpc.state_bytes(), dtype=torch.uint8, device=self.dev)
        self.mpc.state_save(buf)
        return buf

    def restore(self, buf):
        self.mpc.state_load(buf)
