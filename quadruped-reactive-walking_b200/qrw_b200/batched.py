"""Batched solver objects over the C ABI (include/qrw_b200.h).

Arrays are in the ABI layout: instance-major, column-major inside an instance (what
``Eigen::MatrixXd::data()`` holds in the reference).  ``pack_cols`` converts a stack of natural
``[B, rows, cols]`` numpy matrices.  Host inputs (numpy) go through the ``*_host`` entry points
(H2D/D2H inside the call); CUDA tensors (torch) go through the ``*_device`` entry points and stay
asynchronous on the current torch stream.
"""
import ctypes as C

import numpy as np

from ._lib import check, load_library

N_STEPS = 16


def pack_cols(mats):
    """[B, r, c] natural indexing -> [B, r*c] column-major per instance."""
    mats = np.asarray(mats, dtype=np.float64)
    return np.ascontiguousarray(mats.transpose(0, 2, 1)).reshape(mats.shape[0], -1)


def unpack_cols(flat, rows, cols):
    """[B, r*c] column-major per instance -> [B, r, c]."""
    flat = np.asarray(flat)
    return flat.reshape(flat.shape[0], cols, rows).transpose(0, 2, 1)


def _is_cuda_tensor(a):
    return hasattr(a, "is_cuda") and a.is_cuda


def _host(a, n):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if a.size != n:
        raise ValueError(f"expected {n} doubles, got {a.size}")
    return a


def _dev_ptr(t, n):
    import torch

    if t.dtype != torch.float64 or not t.is_contiguous() or t.numel() != n:
        raise ValueError(f"expected a contiguous float64 CUDA tensor with {n} elements")
    return C.c_void_p(t.data_ptr())


def _torch_stream():
    import torch

    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class _DevView:
    """Expose a raw device pointer to torch through __cuda_array_interface__ (no copy)."""

    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f8", "data": (int(ptr), False), "version": 2}


class BatchedMPC:
    """B independent ``MPC`` objects (reference src/MPC.cpp) on one GPU.

    run(num_iter, xref, fsteps) / get_latest_result() / get_gait() / get_Sgait() keep the reference
    names (python/gepadd.cpp:22-31); xref is [B, 12*(n+1)], fsteps [B, N_gait*12], result [B, 24*n].
    """

    def __init__(self, batch, dt=0.02, n_steps=N_STEPS, T_gait=0.32, N_gait=20, device=0, library=None):
        self._L = load_library(library)  # `library`: another build of the same ABI (the A/B test library)
        self.B, self.n, self.N_gait, self.device = int(batch), int(n_steps), int(N_gait), int(device)
        h = C.c_void_p()
        check(self._L.qrw_mpc_create(C.byref(h), self.B, float(dt), self.n, float(T_gait), self.N_gait, self.device))
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            self._L.qrw_mpc_destroy(self._h)
            self._h = None

    __del__ = close

    # -- reference-named API
    def run(self, num_iter, xref, fsteps):
        nx, nf = self.B * 12 * (self.n + 1), self.B * 12 * self.N_gait
        if _is_cuda_tensor(xref):
            check(self._L.qrw_mpc_run_device(self._h, int(num_iter), _dev_ptr(xref, nx), _dev_ptr(fsteps, nf), _torch_stream()))
        else:
            x, f = _host(xref, nx), _host(fsteps, nf)
            check(self._L.qrw_mpc_run_host(self._h, int(num_iter), x.ctypes.data, f.ctypes.data))
        return 0

    def get_latest_result(self, out=None):
        if out is None:
            out = np.empty((self.B, 24 * self.n))
        elif out.dtype != np.float64 or not out.flags.c_contiguous or out.size != self.B * 24 * self.n:
            raise ValueError("out must be a contiguous float64 array of B*24*n elements")
        check(self._L.qrw_mpc_result_host(self._h, out.ctypes.data))
        return out

    def get_gait(self):
        out = np.empty((self.B, 4 * self.N_gait))
        check(self._L.qrw_mpc_gait_host(self._h, out.ctypes.data))
        return out

    def get_Sgait(self):
        out = np.empty((self.B, 12 * self.n))
        check(self._L.qrw_mpc_sgait_host(self._h, out.ctypes.data))
        return out

    # -- batched extras
    def result_tensor(self):
        """Zero-copy torch view [B, 24*n] of the device result buffer."""
        import torch

        p = C.c_void_p()
        check(self._L.qrw_mpc_result_device(self._h, C.byref(p)))
        return torch.as_tensor(_DevView(p.value, (self.B, 24 * self.n)), device=f"cuda:{self.device}")

    def info(self):
        it = np.empty(self.B, np.int32)
        st = np.empty(self.B, np.int32)
        rho = np.empty(self.B)
        nf = np.empty(self.B, np.int32)
        check(self._L.qrw_mpc_info_host(self._h, it.ctypes.data, st.ctypes.data, rho.ctypes.data, nf.ctypes.data))
        return dict(iters=it, status=st, rho=rho, n_factor=nf)

    def set_max_iter(self, k):
        check(self._L.qrw_mpc_set_max_iter(self._h, int(k)))

    def set_kernel(self, variant):
        """2 = the shipped kernel (force-first elimination, twisted state chain, tensor-memory rows).  The A/B test library
        (tests/ab, -DQRW_AB_KERNELS) also has 4 = super-stage state chain, 3 = nested-dissection state solve, 1 = v1 CTA
        kernel, 0 = warp per robot."""
        check(self._L.qrw_mpc_set_kernel(self._h, int(variant)))

    def set_work_order(self, longest_first):
        """True (default): the persistent grid takes the robots with the longest expected solve first; False: index order."""
        check(self._L.qrw_mpc_set_work_order(self._h, int(bool(longest_first))))

    def enable_debug(self):
        check(self._L.qrw_mpc_debug_host(self._h, None))

    def debug(self):
        out = np.empty((self.B, 1281))
        check(self._L.qrw_mpc_debug_host(self._h, out.ctypes.data))
        return dict(D=out[:, :384], E=out[:, 384:1088], c=out[:, 1088], bounds=out[:, 1089:])

    def state_bytes(self):
        n = C.c_size_t()
        check(self._L.qrw_mpc_state_bytes(self._h, C.byref(n)))
        return n.value

    def state_save(self, buf):
        """buf: torch uint8 CUDA tensor of state_bytes()."""
        check(self._L.qrw_mpc_state_save(self._h, C.c_void_p(buf.data_ptr()), _torch_stream()))

    def state_load(self, buf):
        check(self._L.qrw_mpc_state_load(self._h, C.c_void_p(buf.data_ptr()), _torch_stream()))

    def pinned(self):
        """numpy views of the handle's pinned staging buffers (xref, fsteps, result)."""
        px, pf, pr = C.c_void_p(), C.c_void_p(), C.c_void_p()
        check(self._L.qrw_mpc_pinned(self._h, C.byref(px), C.byref(pf), C.byref(pr)))

        def view(p, n):
            return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_double)), shape=(n,))

        return (view(px, self.B * 12 * (self.n + 1)).reshape(self.B, -1), view(pf, self.B * 12 * self.N_gait).reshape(self.B, -1),
                view(pr, self.B * 24 * self.n).reshape(self.B, -1))


class BatchedQPWBC:
    """B independent ``QPWBC`` objects (reference src/QPWBC.cpp).  run(M, Jc, f_cmd, RNEA) with
    M [B, 324], Jc [B, 216] (column-major 18x18 / 12x18), f_cmd [B, 12], RNEA [B, 6]."""

    def __init__(self, batch, device=0):
        self._L = load_library()
        self.B, self.device = int(batch), int(device)
        h = C.c_void_p()
        check(self._L.qrw_wbc_create(C.byref(h), self.B, self.device))
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            self._L.qrw_wbc_destroy(self._h)
            self._h = None

    __del__ = close

    def run(self, M, Jc, f_cmd, RNEA, k_contacts=None):
        B = self.B
        if _is_cuda_tensor(M):
            check(self._L.qrw_wbc_run_device(self._h, _dev_ptr(M, B * 324), _dev_ptr(Jc, B * 216), _dev_ptr(f_cmd, B * 12),
                                             _dev_ptr(RNEA, B * 6), _torch_stream()))
        else:
            a = [_host(M, B * 324), _host(Jc, B * 216), _host(f_cmd, B * 12), _host(RNEA, B * 6)]
            check(self._L.qrw_wbc_run_host(self._h, *[x.ctypes.data for x in a]))
        return 0

    compute = run  # north-star alias

    def results(self, with_H=True):
        f, d = np.empty((self.B, 12)), np.empty((self.B, 6))
        H = np.empty((self.B, 144)) if with_H else None
        check(self._L.qrw_wbc_result_host(self._h, f.ctypes.data, d.ctypes.data, H.ctypes.data if with_H else None))
        return f, d, H

    def get_f_res(self):
        return self.results(False)[0]

    def get_ddq_res(self):
        return self.results(False)[1]

    def get_H(self):
        return self.results(True)[2]

    def info(self):
        it = np.empty(self.B, np.int32)
        st = np.empty(self.B, np.int32)
        rho = np.empty(self.B)
        check(self._L.qrw_wbc_info_host(self._h, it.ctypes.data, st.ctypes.data, rho.ctypes.data))
        return dict(iters=it, status=st, rho=rho)


class BatchedInvKin:
    """Batched ``InvKin::refreshAndCompute`` (reference src/InvKin.cpp:23-73); stateless."""

    def __init__(self, batch, device=0):
        self._L = load_library()
        self.B = int(batch)
        self.ddq = self.dq_cmd = self.q_step = None

    def refreshAndCompute(self, contacts, goals, vgoals, agoals, posf, vf, wf, af, Jf):
        B = self.B
        sizes = [4, 12, 12, 12, 12, 12, 12, 12, 144]
        args = [contacts, goals, vgoals, agoals, posf, vf, wf, af, Jf]
        if _is_cuda_tensor(contacts):
            import torch

            outs = [torch.empty(B, 12, dtype=torch.float64, device=contacts.device) for _ in range(3)]
            check(self._L.qrw_invkin_run_device(B, *[_dev_ptr(a, B * s) for a, s in zip(args, sizes)],
                                                *[C.c_void_p(o.data_ptr()) for o in outs], _torch_stream()))
        else:
            a = [_host(x, B * s) for x, s in zip(args, sizes)]
            outs = [np.empty((B, 12)) for _ in range(3)]
            check(self._L.qrw_invkin_run_host(B, *[x.ctypes.data for x in a], *[o.ctypes.data for o in outs]))
        self.ddq, self.dq_cmd, self.q_step = outs
        return self.ddq

    def get_q_step(self):
        return self.q_step

    def get_dq_cmd(self):
        return self.dq_cmd


class BatchedWbcTick:
    """B robots' ``wbc_controller.compute`` (reference scripts/QP_WBC.py:52-131): Solo12 kinematics / RNEA, InvKin,
    QPWBC and the feed-forward torque in ONE kernel (groups of 8 / 4 / 2 robots per warp, csrc/wbc_tick_core.cuh; Jc / RNEA never
    leave the SM; ``group`` fixes the group size, default: chosen from the batch size).  compute(q, dq, f_cmd, contacts, pgoals, vgoals, agoals) with
    q [B,19], dq [B,18], f_cmd [B,12], contacts [B,4], goals [B,12] (3x4 column-major); results in tau_ff, qdes, vdes,
    f_with_delta, ddq_cmd ([B,12] each)."""

    kernels_per_tick = 1

    def __init__(self, batch, device=0, debug=False, group=0):
        self._L = load_library()
        self.B, self.device = int(batch), int(device)
        h = C.c_void_p()
        check(self._L.qrw_wbctick_create(C.byref(h), self.B, self.device))
        self._h = h
        if group:
            check(self._L.qrw_wbctick_set_group(self._h, int(group)))
        if debug:  # debug(): also copy the intermediates handed to QPWBC (Jc, RNEA) to HBM every tick
            check(self._L.qrw_wbctick_set_debug(self._h, 1))
        self.tau_ff = self.qdes = self.vdes = self.f_with_delta = self.ddq_cmd = None

    def close(self):
        if getattr(self, "_h", None):
            self._L.qrw_wbctick_destroy(self._h)
            self._h = None

    __del__ = close

    def compute(self, q, dq, f_cmd, contacts, pgoals, vgoals, agoals, fetch=True, out=None):
        """out (optional, host path): five [B, 12] float64 arrays that receive tau_ff, qdes, vdes, f_with_delta, ddq_cmd (e.g.
        pinned buffers, so that the device-to-host copies are not staged through pageable memory)."""
        B = self.B
        sizes = [19, 18, 12, 4, 12, 12, 12]
        args = [q, dq, f_cmd, contacts, pgoals, vgoals, agoals]
        if _is_cuda_tensor(q):
            check(self._L.qrw_wbctick_run_device(self._h, *[_dev_ptr(a, B * s) for a, s in zip(args, sizes)], _torch_stream()))
            return 0
        a = [_host(x, B * s) for x, s in zip(args, sizes)]
        check(self._L.qrw_wbctick_run_host(self._h, *[x.ctypes.data for x in a]))
        if fetch:
            outs = [np.empty((B, 12)) for _ in range(5)] if out is None else list(out)
            if len(outs) != 5 or any(o.dtype != np.float64 or not o.flags.c_contiguous or o.size != B * 12 for o in outs):
                raise ValueError("out must be five contiguous float64 arrays of B*12 elements")
            check(self._L.qrw_wbctick_result_host(self._h, *[o.ctypes.data for o in outs]))
            self.tau_ff, self.qdes, self.vdes, self.f_with_delta, self.ddq_cmd = outs
        return 0

    def compute_from_mpc(self, q, dq, mpc_result, contacts, pgoals, vgoals, agoals, n_steps=N_STEPS):
        """compute() with f_cmd = x_f_mpc[12:24, 0] read in place from an MPC result tensor [B, 24 n] (no gather): CUDA tensors."""
        B = self.B
        sizes = [19, 18, 4, 12, 12, 12]
        args = [q, dq, contacts, pgoals, vgoals, agoals]
        p = [_dev_ptr(a, B * s) for a, s in zip(args, sizes)]
        f = C.c_void_p(mpc_result.data_ptr() + 12 * 8)
        check(self._L.qrw_wbctick_run_fstride_device(self._h, p[0], p[1], f, 24 * int(n_steps), p[2], p[3], p[4], p[5], _torch_stream()))
        return 0

    def result_tensors(self):
        """Zero-copy torch views (tau_ff, qdes, vdes, f_res), [B,12] each."""
        import torch

        ps = [C.c_void_p() for _ in range(4)]
        check(self._L.qrw_wbctick_result_device(self._h, *[C.byref(p) for p in ps]))
        return [torch.as_tensor(_DevView(p.value, (self.B, 12)), device=f"cuda:{self.device}") for p in ps]

    def debug(self):
        Jc, r, d = np.empty((self.B, 216)), np.empty((self.B, 6)), np.empty((self.B, 6))
        check(self._L.qrw_wbctick_debug_host(self._h, Jc.ctypes.data, r.ctypes.data, d.ctypes.data))
        return dict(Jc=Jc, RNEA=r, ddq_res=d)


def solo12_feet(q12, dq12):
    """Fixed-base foot kinematics (reference scripts/solo12InvKin.py:47-59) on the GPU: posf, vf, wf, af [B,12], Jf [B,144]."""
    L = load_library()
    q12, dq12 = np.ascontiguousarray(q12, np.float64), np.ascontiguousarray(dq12, np.float64)
    B = q12.size // 12
    outs = [np.empty((B, 12)) for _ in range(4)] + [np.empty((B, 144))]
    check(L.qrw_solo12_feet_host(B, q12.ctypes.data, dq12.ctypes.data, *[o.ctypes.data for o in outs]))
    return outs


def solo12_base_inertia():
    L = load_library()
    Y = np.empty(6)
    check(L.qrw_solo12_base_inertia(Y.ctypes.data))
    return Y
