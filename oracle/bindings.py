"""ctypes bindings of the CPU oracle (oracle/_build/libqrw_oracle.so).

TEST INFRASTRUCTURE ONLY: import this from tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs -- never from the product package.

Every array is instance-major; inside an instance matrices are column-major (Eigen's layout),
e.g. xref is [B, 12*(n+1)] with element (r, c) of instance b at [b, r + 12*c].
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_build", "libqrw_oracle.so")
_REF = os.path.join(_HERE, "_ref", "libst_to_cc.so")

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)


def build(force=False):
    """Compile the oracle (and oracle/_ref when /root/reference is present)."""
    if force or not os.path.exists(_LIB) or any(
        os.path.getmtime(os.path.join(_HERE, f)) > os.path.getmtime(_LIB)
        for f in os.listdir(_HERE)
        if f.endswith((".cpp", ".hpp", ".h"))
    ):
        subprocess.check_call(["make", "-C", _HERE, "-s"], stdout=subprocess.DEVNULL)
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB)
        L.orc_mpc_create.restype = C.c_void_p
        L.orc_mpc_create.argtypes = [C.c_int, C.c_double, C.c_int, C.c_double, C.c_int]
        L.orc_mpc_destroy.argtypes = [C.c_void_p]
        L.orc_mpc_run.argtypes = [C.c_void_p, C.c_int, _dp, _dp, C.c_int]
        L.orc_mpc_result.argtypes = [C.c_void_p, _dp]
        L.orc_mpc_gait.argtypes = [C.c_void_p, _dp]
        L.orc_mpc_sgait.argtypes = [C.c_void_p, _dp]
        L.orc_mpc_info.argtypes = [C.c_void_p, _ip, _ip, _dp, _ip]
        L.orc_mpc_nnz.argtypes = [C.c_void_p, C.c_int]
        L.orc_mpc_ml.argtypes = [C.c_void_p, C.c_int, _dp, _ip, _ip]
        L.orc_mpc_bounds.argtypes = [C.c_void_p, C.c_int, _dp, _dp]
        L.orc_mpc_solver_state.argtypes = [C.c_void_p, C.c_int] + [_dp] * 10
        L.orc_mpc_set_trace.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.orc_mpc_trace.argtypes = [C.c_void_p, C.c_int, _dp, C.c_int]
        L.orc_mpc_set_max_iter.argtypes = [C.c_void_p, C.c_int]
        L.orc_wbc_create.restype = C.c_void_p
        L.orc_wbc_create.argtypes = [C.c_int]
        L.orc_wbc_destroy.argtypes = [C.c_void_p]
        L.orc_wbc_run.argtypes = [C.c_void_p, _dp, _dp, _dp, _dp, C.c_int]
        L.orc_wbc_result.argtypes = [C.c_void_p, _dp, _dp, _dp]
        L.orc_wbc_info.argtypes = [C.c_void_p, _ip, _ip, _dp]
        L.orc_pinv6.argtypes = [_dp, _dp]
        L.orc_invkin_run.argtypes = [C.c_int] + [_dp] * 12
        L.orc_osqp_dense.argtypes = [C.c_int, C.c_int] + [_dp] * 5 + [C.c_double, C.c_int, _dp, _dp, _ip]
        L.orc_osqp_dense_ex.argtypes = ([C.c_int, C.c_int] + [_dp] * 5 + [C.c_double, C.c_double, C.c_int, C.c_int, _dp, _dp, _ip, _dp])
        L.orc_solo12_feet.argtypes = [C.c_int] + [_dp] * 7
        L.orc_solo12_rnea.argtypes = [C.c_int] + [_dp] * 4
        L.orc_solo12_jc.argtypes = [C.c_int] + [_dp] * 3
        L.orc_solo12_crba.argtypes = [_dp, _dp]
        L.orc_planner_create.restype = C.c_void_p
        L.orc_planner_create.argtypes = [C.c_int] + [C.c_double] * 4 + [C.c_int, C.c_int, C.c_double, _dp, _dp, _dp,
                                                                         C.c_double, C.c_double]
        L.orc_planner_destroy.argtypes = [C.c_void_p]
        L.orc_planner_run.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp, _dp, _dp, _ip]
        L.orc_planner_outputs.argtypes = [C.c_void_p] + [_dp] * 7
        L.orc_planner_state.argtypes = [C.c_void_p] + [_dp] * 5 + [_ip]
        L.orc_planner_phase_duration.restype = C.c_double
        L.orc_planner_phase_duration.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_double, _dp]
        L.orc_planner_create_walk.argtypes = [C.c_void_p, C.c_int]
        _lib = L
    return _lib


def _d(a):
    return a.ctypes.data_as(_dp)


def _i(a):
    return a.ctypes.data_as(_ip)


def _c(a):
    return np.ascontiguousarray(a, dtype=np.float64)


class OracleMPC:
    """Batch of independent restated `MPC` objects (reference src/MPC.cpp)."""

    def __init__(self, batch=1, dt=0.02, n_steps=16, T_gait=0.32, N_gait=20):
        self.B, self.n, self.N_gait = batch, n_steps, N_gait
        self.h = lib().orc_mpc_create(batch, dt, n_steps, T_gait, N_gait)

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_mpc_destroy(self.h)
            self.h = None

    def run(self, num_iter, xref, fsteps, nthreads=1):
        xref, fsteps = _c(xref), _c(fsteps)
        assert xref.size == self.B * 12 * (self.n + 1) and fsteps.size == self.B * 12 * self.N_gait
        lib().orc_mpc_run(self.h, int(num_iter), _d(xref), _d(fsteps), int(nthreads))
        return 0

    def get_latest_result(self):
        out = np.empty((self.B, 24 * self.n))
        lib().orc_mpc_result(self.h, _d(out))
        return out

    def get_gait(self):
        out = np.empty((self.B, 4 * self.N_gait))
        lib().orc_mpc_gait(self.h, _d(out))
        return out

    def get_Sgait(self):
        out = np.empty((self.B, 12 * self.n))
        lib().orc_mpc_sgait(self.h, _d(out))
        return out

    def info(self):
        it = np.empty(self.B, np.int32)
        stt = np.empty(self.B, np.int32)
        rho = np.empty(self.B)
        nf = np.empty(self.B, np.int32)
        lib().orc_mpc_info(self.h, _i(it), _i(stt), _d(rho), _i(nf))
        return dict(iters=it, status=stt, rho=rho, n_factor=nf)

    def ml(self, b=0):
        nnz = lib().orc_mpc_nnz(self.h, b)
        x = np.empty(nnz)
        ri = np.empty(nnz, np.int32)
        cp = np.empty(24 * self.n + 1, np.int32)
        lib().orc_mpc_ml(self.h, b, _d(x), _i(ri), _i(cp))
        return x, ri, cp

    def bounds(self, b=0):
        l = np.empty(44 * self.n)
        u = np.empty(44 * self.n)
        lib().orc_mpc_bounds(self.h, b, _d(l), _d(u))
        return l, u

    def solver_state(self, b=0):
        n, m = 24 * self.n, 44 * self.n
        nnz = lib().orc_mpc_nnz(self.h, b)
        D, E, c = np.empty(n), np.empty(m), np.empty(1)
        x, z, y = np.empty(n), np.empty(m), np.empty(m)
        As, Ps, ls, us = np.empty(nnz), np.empty(n), np.empty(m), np.empty(m)
        lib().orc_mpc_solver_state(self.h, b, _d(D), _d(E), _d(c), _d(x), _d(z), _d(y), _d(As), _d(Ps), _d(ls), _d(us))
        return dict(D=D, E=E, c=c[0], x=x, z=z, y=y, A=As, P=Ps, l=ls, u=us)

    def set_trace(self, b=0, on=True):
        lib().orc_mpc_set_trace(self.h, b, int(on))

    def trace(self, b=0):
        nrow = lib().orc_mpc_trace(self.h, b, None, 0)
        out = np.empty((nrow, 4))
        lib().orc_mpc_trace(self.h, b, _d(out), nrow)
        return out

    def set_max_iter(self, k):
        lib().orc_mpc_set_max_iter(self.h, int(k))


class OracleWBC:
    """Batch of independent restated `QPWBC` objects (reference src/QPWBC.cpp)."""

    def __init__(self, batch=1):
        self.B = batch
        self.h = lib().orc_wbc_create(batch)

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_wbc_destroy(self.h)
            self.h = None

    def run(self, M, Jc, f_cmd, RNEA, nthreads=1):
        M, Jc, f_cmd, RNEA = _c(M), _c(Jc), _c(f_cmd), _c(RNEA)
        assert M.size == self.B * 324 and Jc.size == self.B * 216
        lib().orc_wbc_run(self.h, _d(M), _d(Jc), _d(f_cmd), _d(RNEA), int(nthreads))
        return 0

    def results(self):
        f, d, H = np.empty((self.B, 12)), np.empty((self.B, 6)), np.empty((self.B, 144))
        lib().orc_wbc_result(self.h, _d(f), _d(d), _d(H))
        return f, d, H

    def info(self):
        it = np.empty(self.B, np.int32)
        stt = np.empty(self.B, np.int32)
        rho = np.empty(self.B)
        lib().orc_wbc_info(self.h, _i(it), _i(stt), _d(rho))
        return dict(iters=it, status=stt, rho=rho)


def invkin(contacts, goals, vgoals, agoals, posf, vf, wf, af, Jf):
    """Batched restated InvKin::refreshAndCompute (reference src/InvKin.cpp:23-73)."""
    B = _c(contacts).size // 4
    args = [_c(a) for a in (contacts, goals, vgoals, agoals, posf, vf, wf, af, Jf)]
    ddq, dq, qs = np.empty((B, 12)), np.empty((B, 12)), np.empty((B, 12))
    lib().orc_invkin_run(B, *[_d(a) for a in args], _d(ddq), _d(dq), _d(qs))
    return ddq, dq, qs


def pinv6(a):
    a = np.asfortranarray(a, dtype=np.float64)
    out = np.empty((6, 6), order="F")
    lib().orc_pinv6(_d(a), _d(out))
    return out


def osqp_dense(P, A, q, l, u, eps=1e-8, max_iter=20000):
    """Restated OSQP on a dense problem; returns (x, y, iters, status)."""
    P = np.asfortranarray(P, dtype=np.float64)
    A = np.asfortranarray(A, dtype=np.float64)
    m, n = A.shape
    q, l, u = _c(q), _c(l), _c(u)
    x, y = np.empty(n), np.empty(m)
    it = np.zeros(1, np.int32)
    st = lib().orc_osqp_dense(n, m, _d(P), _d(A), _d(q), _d(l), _d(u), float(eps), int(max_iter), _d(x), _d(y), _i(it))
    return x, y, int(it[0]), st


def osqp_dense_ex(P, A, q, l, u, eps=1e-3, alpha=1.6, max_iter=4000, adaptive_rho_interval=0):
    """Restated OSQP with its DEFAULT settings except those given; returns (x, y, iters, status, info) with the last iterate
    (also when max_iter stops the run) and info = dict(pri_res, dua_res, rho, rho_updates, rho_estimate): what OSQP prints."""
    P = np.asfortranarray(P, dtype=np.float64)
    A = np.asfortranarray(A, dtype=np.float64)
    m, n = A.shape
    q, l, u = _c(q), _c(l), _c(u)
    x, y, info = np.empty(n), np.empty(m), np.empty(5)
    it = np.zeros(1, np.int32)
    st = lib().orc_osqp_dense_ex(n, m, _d(P), _d(A), _d(q), _d(l), _d(u), float(eps), float(alpha), int(max_iter),
                                 int(adaptive_rho_interval), _d(x), _d(y), _i(it), _d(info))
    return x, y, int(it[0]), st, dict(pri_res=info[0], dua_res=info[1], rho=info[2], rho_updates=int(info[3]), rho_estimate=info[4])


# ---------------------------------------------------------------------------------------------
# Solo12 rigid-body quantities (what the reference gets from Pinocchio; oracle/solo12_oracle.cpp)
# ---------------------------------------------------------------------------------------------
def solo12_feet(q12, dq12):
    """solo12InvKin.py:47-59: (posf, vf, wf, af [B,12] as 4x3 column-major, Jf [B,144] as 12x12 column-major)."""
    q12, dq12 = _c(q12), _c(dq12)
    B = q12.size // 12
    outs = [np.empty((B, 12)) for _ in range(4)] + [np.empty((B, 144))]
    lib().orc_solo12_feet(B, _d(q12), _d(dq12), *[_d(o) for o in outs])
    return outs


def solo12_rnea(q, v, a):
    q, v, a = _c(q), _c(v), _c(a)
    B = q.size // 19
    tau = np.empty((B, 18))
    lib().orc_solo12_rnea(B, _d(q), _d(v), _d(a), _d(tau))
    return tau


def solo12_jc(q, contacts):
    q, contacts = _c(q), _c(contacts)
    B = q.size // 19
    Jc = np.empty((B, 216))
    lib().orc_solo12_jc(B, _d(q), _d(contacts), _d(Jc))
    return Jc


def solo12_crba(q):
    q = _c(q)
    M = np.empty((18, 18), order="F")
    lib().orc_solo12_crba(_d(q), _d(M))
    return M


def wbc_tick(owbc, q, dq, f_cmd, contacts, pgoals, vgoals, agoals, nthreads=1):
    """wbc_controller.compute (scripts/QP_WBC.py:52-131) for a batch, composed from the oracle's pieces.
    owbc: OracleWBC carrying the QP's warm start.  Returns dict(tau_ff, qdes, vdes, f_res, ddq_cmd, Jc, RNEA, ddq_res)."""
    q, dq = _c(q).reshape(-1, 19), _c(dq).reshape(-1, 18)
    B = q.shape[0]
    posf, vf, wf, af, Jf = solo12_feet(q[:, 7:], dq[:, 6:])
    ddq, dq_cmd, q_step = invkin(contacts, pgoals, vgoals, agoals, posf, vf, wf, af, Jf)
    q0 = np.zeros(19)
    q0[6] = 1.0
    M = solo12_crba(q0)
    M[:6, :6] = M[:6, :6] * np.eye(6)  # QP_WBC.py:93
    Mb = np.tile(np.asfortranarray(M).ravel(order="F"), (B, 1))
    Jc = solo12_jc(q, contacts)
    acc = np.zeros((B, 18))
    acc[:, 6:] = ddq
    RNEA = solo12_rnea(q, dq, acc)[:, :6]
    owbc.run(Mb, Jc, f_cmd, RNEA, nthreads)
    f_res, ddq_res, _ = owbc.results()
    acc[:, :6] += ddq_res
    tau = solo12_rnea(q, dq, acc)[:, 6:]
    Jm = Jc.reshape(B, 18, 12).transpose(0, 2, 1)  # [B, 12, 18]
    tau_ff = tau - np.einsum("bij,bi->bj", Jm[:, :, 6:], f_res)
    return dict(tau_ff=tau_ff, qdes=q[:, 7:] + q_step, vdes=dq_cmd, f_res=f_res, ddq_cmd=ddq, Jc=Jc, RNEA=RNEA,
                ddq_res=ddq_res)


# ---------------------------------------------------------------------------------------------
# Reference-built converter (oracle/_ref): the reference's own src/st_to_cc.cpp compiled in place.
# ---------------------------------------------------------------------------------------------
def ref_st_to_cc_available():
    return os.path.exists(_REF)


def ref_st_to_cc(rows, cols, vals, n_cols):
    """Run the REFERENCE st_to_cc_size/index/values on a triplet list; returns (icc, ccc, acc)."""
    L = C.CDLL(_REF)
    # C++ symbols (the reference builds this file as C++): look them up by mangled name
    size = getattr(L, "_Z13st_to_cc_sizeiPiS_")
    index = getattr(L, "_Z14st_to_cc_indexiPiS_iiS_S_")
    values = getattr(L, "_Z15st_to_cc_valuesiPiS_PdiiS_S_")
    rows = np.ascontiguousarray(rows, np.int32)
    cols = np.ascontiguousarray(cols, np.int32)
    vals = _c(vals)
    nst = rows.size
    size.restype = C.c_int
    ncc = size(nst, _i(rows), _i(cols))
    icc = np.zeros(ncc, np.int32)
    ccc = np.zeros(n_cols + 1, np.int32)
    index(nst, _i(rows), _i(cols), ncc, n_cols, _i(icc), _i(ccc))
    values.restype = _dp
    p = values(nst, _i(rows), _i(cols), _d(vals), ncc, n_cols, _i(icc), _i(ccc))
    acc = np.array([p[k] for k in range(ncc)])
    return icc, ccc, acc


SHOULDERS34 = np.array([0.1946, 0.14695, 0.0, 0.1946, -0.14695, 0.0, -0.1946, 0.14695, 0.0, -0.1946, -0.14695, 0.0])


class OraclePlanner:
    """Batch of restated Gait + StatePlanner + FootstepPlanner + FootTrajectoryGenerator objects wired as
    scripts/Controller.py:124-139 (oracle/planner_oracle.cpp).  run(k, ...) = Controller.py:222-236 for tick k."""

    GAIT, FOOTSTEPS, FOOTTRAJ, STATE = 1, 2, 4, 8

    def __init__(self, batch, dt_mpc=0.02, dt_wbc=0.002, T_gait=0.32, T_mpc=0.32, N_gait=20, k_mpc=10, h_ref=0.2447495,
                 shoulders=None, ftg_target=None, ftg_pos=None, max_height=0.05, lock_time=0.07):
        sh = _c(SHOULDERS34 if shoulders is None else shoulders).ravel()
        tg = _c(sh if ftg_target is None else ftg_target).ravel()
        ps = _c(sh if ftg_pos is None else ftg_pos).ravel()
        self.B, self.N_gait, self.n = batch, N_gait, int(round(T_mpc / dt_mpc))
        self.h = lib().orc_planner_create(batch, dt_mpc, dt_wbc, T_gait, T_mpc, N_gait, k_mpc, h_ref, _d(sh), _d(tg), _d(ps),
                                          max_height, lock_time)
        if not self.h:
            raise ValueError("Sizes of matrices are too small for considered durations. Increase N_gait in config file.")

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_planner_destroy(self.h)
            self.h = None

    def run(self, k, q7, v6, vref6, code=None, stages=15):
        q7, v6, vref6 = _c(q7), _c(v6), _c(vref6)
        assert q7.size == 7 * self.B and v6.size == 6 * self.B and vref6.size == 6 * self.B
        cd = None if code is None else np.ascontiguousarray(code, np.int32)
        lib().orc_planner_run(self.h, int(k), int(stages), _d(q7), _d(v6), _d(vref6), None if cd is None else _i(cd))

    def outputs(self):
        B, N = self.B, self.N_gait
        o = dict(xref=np.empty((B, 12 * (self.n + 1))), fsteps=np.empty((B, 12 * N)), gait=np.empty((B, 4 * N)),
                 o_target=np.empty((B, 12)), fpos=np.empty((B, 12)), fvel=np.empty((B, 12)), facc=np.empty((B, 12)))
        lib().orc_planner_outputs(self.h, *[_d(o[k]) for k in ("xref", "fsteps", "gait", "o_target", "fpos", "fvel", "facc")])
        return o

    def state(self):
        B, N = self.B, self.N_gait
        s = dict(past=np.empty((B, 4 * N)), desired=np.empty((B, 4 * N)), cur_fs=np.empty((B, 12)), t0s=np.empty((B, 4)),
                 t_swing=np.empty((B, 4)))
        fl = np.empty((B, 2), np.int32)
        lib().orc_planner_state(self.h, *[_d(s[k]) for k in ("past", "desired", "cur_fs", "t0s", "t_swing")], _i(fl))
        s["flags"] = fl
        return s

    def phase_duration(self, b, i, j, value):
        rem = np.zeros(1)
        t = lib().orc_planner_phase_duration(self.h, b, i, j, float(value), _d(rem))
        return t, rem[0]


def feet_commands(pos, vel, acc, q7, vref6, p_prev, v_prev, h_ref):
    """Controller.py:286-296 for a batch: feet position / velocity / acceleration commands in the base frame from the
    FootTrajectoryGenerator outputs ([B,12], 3x4 column-major), with oRh = Rz(yaw(q)), oTh = (x, y, 0) (:419-423)."""
    B = pos.shape[0]
    x, y, z, w = q7[:, 3], q7[:, 4], q7[:, 5], q7[:, 6]
    yaw = np.arctan2(2 * (x * y + z * w), 1 - 2 * (y * y + z * z))
    c, s = np.cos(yaw), np.sin(yaw)
    RT = np.zeros((B, 3, 3))
    RT[:, 0, 0], RT[:, 0, 1], RT[:, 1, 0], RT[:, 1, 1], RT[:, 2, 2] = c, s, -s, c, 1.0
    P, V, A = (a.reshape(B, 4, 3) for a in (pos, vel, acc))
    pp, vp = p_prev.reshape(B, 4, 3), v_prev.reshape(B, 4, 3)
    wv = np.broadcast_to(vref6[:, None, 3:6], (B, 4, 3))
    a_cmd = np.einsum("bij,bfj->bfi", RT, A) - np.cross(wv, np.cross(wv, pp)) - 2 * np.cross(wv, vp)
    v_cmd = np.einsum("bij,bfj->bfi", RT, V) - vref6[:, None, 0:3] - np.cross(wv, pp)
    off = np.zeros((B, 1, 3))
    off[:, 0, 0], off[:, 0, 1], off[:, 0, 2] = q7[:, 0], q7[:, 1], h_ref
    p_cmd = np.einsum("bij,bfj->bfi", RT, P - off)
    return p_cmd.reshape(B, 12), v_cmd.reshape(B, 12), a_cmd.reshape(B, 12)
