"""Synthetic inputs for the BASELINE.json configs.

Host-side numpy only.  The generators restate, in vectorised form, what the reference's input
producers hand to `MPC.run`:

* reference states `xref` (12 x (n+1)) -- StatePlanner::computeReferenceStates,
  /root/reference/src/StatePlanner.cpp:21-60
* contact schedules -- Gait::create_trot / create_walk / create_pacing / create_bounding /
  create_static, /root/reference/src/Gait.cpp:38-108, rolled one row per MPC tick (Gait.cpp:221-260)
* footstep matrix `fsteps` (N_gait x 12): one (x, y, 0) foothold per foot in contact, exactly 0.0
  for swing feet (that is how MPC::construct_gait detects them, src/MPC.cpp:686-701); footholds
  are the Solo12 shoulders (scripts/Controller.py:131-133) plus uniform noise.

Layouts are the batched C-ABI's: instance-major, column-major inside an instance
(`xref[b, r + 12*c]`, `fsteps[b, k + N_gait*j]`).
"""
import numpy as np

DT_MPC = 0.02
N_STEPS = 16
T_GAIT = 0.32
N_GAIT = 20
H_REF = 0.2027682065963745  # (double)0.2027682f, src/MPC.cpp:27-28
Z_INIT = 0.24474949993103629  # scripts/test_mpc.py:39
SHOULDERS = np.array([[0.1946, 0.1946, -0.1946, -0.1946], [0.14695, -0.14695, 0.14695, -0.14695]])

GAITS = ("trot", "walk", "bounding", "pacing", "static")

# Algorithmic FP64 work of the block-tridiagonal ADMM (DESIGN.md "FLOP model"): per iteration 16 dense
# 24x24 mat-vecs (forward) + 15 dense 24x12 (backward) + two passes over the 1998-nnz constraint
# matrix + ~10 vector ops on 704 rows and 6 on 384 variables; per call 16 Gauss-Jordan 24^3
# inversions with their block assembly and 10 Ruiz passes.
FLOP_PER_ADMM_ITER = 16 * 2 * 576 + 15 * 2 * 288 + 4 * 1998 + 10 * 704 + 6 * 384
FLOP_PER_SETUP = 16 * (2 * 24 ** 3 + 12000) + 10 * 12000


def gait_cycle(name, n=N_STEPS):
    """One period (n rows x 4) of the named gait, rows of 0/1 (Gait.cpp:38-108)."""
    g = np.zeros((n, 4))
    h, q = n // 2, n // 4
    if name == "trot":
        g[:h] = [1, 0, 0, 1]
        g[h:] = [0, 1, 1, 0]
    elif name == "pacing":
        g[:h] = [1, 0, 1, 0]
        g[h:] = [0, 1, 0, 1]
    elif name == "bounding":
        g[:h] = [1, 1, 0, 0]
        g[h:] = [0, 0, 1, 1]
    elif name == "walk":
        g[:] = 1
        for i in range(4):
            g[i * q:(i + 1) * q, i] = 0
    elif name == "static":
        g[:] = 1
    else:
        raise ValueError(name)
    return g


def reference_states(x0, vref, n=N_STEPS, dt=DT_MPC, h_ref=H_REF, z_average=0.0):
    """Vectorised StatePlanner::computeReferenceStates.  x0: [B,12] current state in the
    horizontal frame (x = y = yaw = 0 are imposed as the reference does), vref: [B,6].
    Returns xref [B, 12, n+1]."""
    B = x0.shape[0]
    xr = np.zeros((B, 12, n + 1))
    xr[:, 2, 0] = x0[:, 2]
    xr[:, 3:5, 0] = x0[:, 3:5]
    xr[:, 6:12, 0] = x0[:, 6:12]
    t = np.linspace(dt, n * dt, n)[None, :]  # dt_vector_
    vx, vy, wz = vref[:, 0:1], vref[:, 1:2], vref[:, 5:6]
    nz = wz != 0
    wz_safe = np.where(nz, wz, 1.0)
    xr[:, 0, 1:] = np.where(nz, (vx * np.sin(wz * t) + vy * (np.cos(wz * t) - 1.0)) / wz_safe, vx * t)
    xr[:, 1, 1:] = np.where(nz, (vy * np.sin(wz * t) - vx * (np.cos(wz * t) - 1.0)) / wz_safe, vy * t)
    xr[:, 2, 1:] = h_ref + z_average
    yaw = wz * t
    xr[:, 5, 1:] = yaw
    xr[:, 6, 1:] = vx * np.cos(yaw) - vy * np.sin(yaw)
    xr[:, 7, 1:] = vx * np.sin(yaw) + vy * np.cos(yaw)
    xr[:, 11, 1:] = wz
    return xr


class MpcWorkload:
    """A batch of B robots with a gait timeline long enough for `ticks` rolled MPC calls.

    inputs(t, x0) gives the (xref, fsteps) of MPC tick t with the current state x0 ([B,12], or None
    for the initial random state) in the C-ABI layout.
    """

    def __init__(self, B, seed, gaits=("trot",), ticks=9, n=N_STEPS, N_gait=N_GAIT):
        rng = np.random.default_rng(seed)
        self.B, self.n, self.N_gait, self.ticks = B, n, N_gait, ticks
        # random initial state (ranges of scripts/test_mpc.py:93) and velocity reference
        # (envelope of scripts/Joystick.py:207-278)
        x0 = np.zeros((B, 12))
        x0[:, 2] = rng.uniform(0.18, 0.26, B)
        x0[:, 3:5] = rng.uniform(-0.1, 0.1, (B, 2))
        x0[:, 6:9] = rng.uniform(-0.3, 0.3, (B, 3))
        x0[:, 9:12] = rng.uniform(-0.4, 0.4, (B, 3))
        self.x0 = x0
        vref = np.zeros((B, 6))
        vref[:, 0] = rng.uniform(-0.5, 1.5, B)
        vref[:, 1] = rng.uniform(-0.5, 0.5, B)
        vref[:, 5] = rng.uniform(-0.7, 0.7, B)
        self.vref = vref
        # gait timeline: rows t .. t+n-1 are the horizon of tick t
        L = ticks + n
        self.kind = rng.integers(0, len(gaits), B)
        phase = rng.integers(0, n, B)
        cycles = np.stack([gait_cycle(g, n) for g in gaits])  # [G, n, 4]
        rows = (phase[:, None] + np.arange(L)[None, :]) % n  # [B, L]
        self.timeline = cycles[self.kind[:, None], rows]  # [B, L, 4]
        # one foothold per contiguous stance segment of each foot
        noise = rng.uniform(-0.05, 0.05, (B, L, 4, 2))
        start = np.ones((B, L, 4), bool)
        start[:, 1:] = self.timeline[:, 1:] != self.timeline[:, :-1]
        seg_id = np.cumsum(start, axis=1) - 1  # [B, L, 4]
        seg_noise = np.take_along_axis(noise, np.broadcast_to(seg_id[..., None], noise.shape), axis=1)
        # index trick above picks noise[b, seg_id, foot]: constant inside a segment
        pos = np.zeros((B, L, 4, 3))
        pos[..., 0] = SHOULDERS[0][None, None, :] + seg_noise[..., 0]
        pos[..., 1] = SHOULDERS[1][None, None, :] + seg_noise[..., 1]
        self.footholds = pos * self.timeline[..., None]  # swing feet exactly 0.0

    def gait(self, t):
        """gait matrix the reference planner would hold at tick t: [B, N_gait, 4]."""
        g = np.zeros((self.B, self.N_gait, 4))
        g[:, : self.n] = self.timeline[:, t:t + self.n]
        return g

    def inputs(self, t, x0=None):
        x0 = self.x0 if x0 is None else x0
        xr = reference_states(x0, self.vref, self.n)
        # closed loop as scripts/test_mpc.py:78 -- the whole predicted state becomes column 0
        xr[:, :, 0] = x0
        fs = np.zeros((self.B, self.N_gait, 12))
        fs[:, : self.n] = self.footholds[:, t:t + self.n].reshape(self.B, self.n, 12)
        xref = np.ascontiguousarray(xr.transpose(0, 2, 1)).reshape(self.B, 12 * (self.n + 1))
        fsteps = np.ascontiguousarray(fs.transpose(0, 2, 1)).reshape(self.B, 12 * self.N_gait)
        return xref, fsteps


def config1_single_trot():
    """BASELINE config 1: the single trot solve of scripts/test_mpc.py:136-146 (current signature:
    12-column fsteps, no duration column).  Returns (xref [1,204], fsteps [1,240])."""
    xr = np.zeros((12, N_STEPS + 1))
    xr[2, :] = Z_INIT
    fs = np.zeros((N_GAIT, 12))
    pair_1 = np.array([0.195, 0.147, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, -0.195, -0.147, 0.0])
    pair_2 = np.array([0.0, 0.0, 0.0, 0.195, -0.147, 0.0, -0.195, 0.147, 0.0, 0.0, 0.0, 0.0])
    fs[0:8] = pair_1
    fs[8:16] = pair_2
    return xr.T.reshape(1, -1).copy(), fs.T.reshape(1, -1).copy()


def four_stance(x0=None):
    """scripts/test_mpc.py:54-58 -- four feet on the ground for the whole horizon."""
    xr = np.zeros((12, N_STEPS + 1))
    xr[2, :] = Z_INIT
    if x0 is not None:
        xr[:, 0] = x0
    fs = np.zeros((N_GAIT, 12))
    fs[0:16] = np.array([0.195, 0.147, 0.0, 0.195, -0.147, 0.0, -0.195, 0.147, 0.0, -0.195, -0.147, 0.0])
    return xr.T.reshape(1, -1).copy(), fs.T.reshape(1, -1).copy()


def next_state(result, n=N_STEPS):
    """Predicted next state x_f_applied[:12, 0] of every instance: result is [B, 24*n]."""
    return result[:, :12].copy()


# ---------------------------------------------------------------------------------------------
# QPWBC / InvKin drop-in API inputs (matrices as scripts/QP_WBC.py:88-107 and
# scripts/solo12InvKin.py:47-62 hand them to the C++ classes), synthetic.
# ---------------------------------------------------------------------------------------------
MASS = 2.500002861022949  # (double)2.50000279f, src/MPC.cpp:17
Y_DIAG = np.array([MASS, MASS, MASS, 0.0309249, 0.051061, 0.0693976])  # composite body at q = 0


def _skew(v):
    z = np.zeros(v.shape[:-1])
    return np.stack([np.stack([z, -v[..., 2], v[..., 1]], -1), np.stack([v[..., 2], z, -v[..., 0]], -1),
                     np.stack([-v[..., 1], v[..., 0], z], -1)], -2)


def wbc_inputs(B, seed, swing_prob=0.35):
    """(M [B,324], Jc [B,216], f_cmd [B,12], RNEA [B,6]) in the C-ABI layout.

    M: 18x18 with the 6x6 base block masked to its diagonal (scripts/QP_WBC.py:89-93); Jc rows of a
    stance foot = [I3 | -skew(p_foot) | leg block], zero for swing feet (:96-101); f_cmd = weight
    shared by the stance feet plus noise inside the friction cone, 0 for swing feet; RNEA = base
    wrench of a slowly moving trunk.
    """
    rng = np.random.default_rng(seed)
    M = np.zeros((B, 18, 18))
    M[:, np.arange(6), np.arange(6)] = Y_DIAG
    L = rng.uniform(-0.02, 0.02, (B, 12, 12))
    M[:, 6:, 6:] = np.einsum("bij,bkj->bik", L, L) + 0.01 * np.eye(12)
    M[:, :6, 6:] = rng.uniform(-0.02, 0.02, (B, 6, 12))
    M[:, 6:, :6] = M[:, :6, 6:].transpose(0, 2, 1)
    contact = rng.uniform(size=(B, 4)) > swing_prob
    contact[contact.sum(1) == 0, 0] = True
    p = np.zeros((B, 4, 3))
    p[..., 0] = SHOULDERS[0] + rng.uniform(-0.05, 0.05, (B, 4))
    p[..., 1] = SHOULDERS[1] + rng.uniform(-0.05, 0.05, (B, 4))
    p[..., 2] = -rng.uniform(0.18, 0.26, (B, 1))
    Jc = np.zeros((B, 12, 18))
    for i in range(4):
        Jc[:, 3 * i:3 * i + 3, 0:3] = np.eye(3)
        Jc[:, 3 * i:3 * i + 3, 3:6] = -_skew(p[:, i])
        Jc[:, 3 * i:3 * i + 3, 6 + 3 * i:9 + 3 * i] = rng.uniform(-0.2, 0.2, (B, 3, 3))
    Jc *= np.repeat(contact, 3, axis=1)[:, :, None]
    nc = contact.sum(1, keepdims=True)
    f = np.zeros((B, 4, 3))
    f[..., 2] = MASS * 9.81 / nc + rng.uniform(-1.0, 1.0, (B, 4))
    f[..., 0:2] = rng.uniform(-0.3, 0.3, (B, 4, 2)) * f[..., 2:3]
    f *= contact[..., None]
    RNEA = np.zeros((B, 6))
    RNEA[:, 2] = MASS * 9.81
    RNEA += rng.uniform(-1.0, 1.0, (B, 6)) * np.array([2.0, 2.0, 2.0, 0.3, 0.3, 0.3])
    return pack_cols(M), pack_cols(Jc), f.reshape(B, 12), RNEA


def pack_cols(mats):
    mats = np.asarray(mats, dtype=np.float64)
    return np.ascontiguousarray(mats.transpose(0, 2, 1)).reshape(mats.shape[0], -1)


def invkin_inputs(B, seed):
    """(contacts [B,4], goals, vgoals, agoals [B,12] (3x4 col-major), posf, vf, wf, af [B,12]
    (4x3 col-major), Jf [B,144] (12x12 col-major)) for InvKin::refreshAndCompute."""
    rng = np.random.default_rng(seed)
    contacts = (rng.uniform(size=(B, 4)) > 0.4).astype(np.float64)
    posf = np.zeros((B, 4, 3))
    posf[..., 0] = SHOULDERS[0] + rng.uniform(-0.03, 0.03, (B, 4))
    posf[..., 1] = SHOULDERS[1] + rng.uniform(-0.03, 0.03, (B, 4))
    posf[..., 2] = -0.22 + rng.uniform(-0.03, 0.03, (B, 4))
    goals = posf.transpose(0, 2, 1) + rng.uniform(-0.03, 0.03, (B, 3, 4))
    vgoals = rng.uniform(-0.5, 0.5, (B, 3, 4))
    agoals = rng.uniform(-5.0, 5.0, (B, 3, 4))
    vf = rng.uniform(-0.5, 0.5, (B, 4, 3))
    wf = rng.uniform(-2.0, 2.0, (B, 4, 3))
    af = rng.uniform(-3.0, 3.0, (B, 4, 3))
    Jf = rng.uniform(-0.05, 0.05, (B, 12, 12))
    for i in range(4):  # well-conditioned diagonal 3x3 leg blocks
        Jf[:, 3 * i:3 * i + 3, 3 * i:3 * i + 3] += 0.16 * np.eye(3) + rng.uniform(-0.08, 0.08, (B, 3, 3))
    return (contacts, pack_cols(goals), pack_cols(vgoals), pack_cols(agoals), pack_cols(posf), pack_cols(vf),
            pack_cols(wf), pack_cols(af), pack_cols(Jf))


Q_INIT = np.array([0.0, 0.7, -1.4, 0.0, 0.7, -1.4, 0.0, -0.7, 1.4, 0.0, -0.7, 1.4])  # scripts/main_solo12_control.py q_init


def wbc_tick_inputs(B, seed, level_base=True):
    """Inputs of wbc_controller.compute (scripts/QP_WBC.py:52) for B robots, BASELINE config 4:
    q [B,19] (base at h_ref; identity orientation as scripts/Controller.py:275-283 does, or a small random tilt),
    dq [B,18], f_cmd [B,12], contacts [B,4], pgoals / vgoals / agoals [B,12] (3x4 column-major)."""
    rng = np.random.default_rng(seed)
    q = np.zeros((B, 19))
    q[:, 2] = H_REF
    q[:, 6] = 1.0
    if not level_base:
        ax = rng.normal(size=(B, 3))
        ax /= np.linalg.norm(ax, axis=1, keepdims=True)
        ang = rng.uniform(-0.2, 0.2, B)
        q[:, 3:6] = ax * np.sin(ang / 2)[:, None]
        q[:, 6] = np.cos(ang / 2)
    q[:, 7:] = Q_INIT + rng.uniform(-0.3, 0.3, (B, 12))
    dq = np.zeros((B, 18))
    dq[:, 0:3] = rng.uniform(-0.5, 0.5, (B, 3))
    dq[:, 3:6] = rng.uniform(-0.5, 0.5, (B, 3))
    dq[:, 6:] = rng.uniform(-3.0, 3.0, (B, 12))
    contacts = (rng.uniform(size=(B, 4)) > 0.35).astype(np.float64)
    contacts[contacts.sum(1) == 0, 0] = 1.0
    nc = contacts.sum(1, keepdims=True)
    f = np.zeros((B, 4, 3))
    f[..., 2] = MASS * 9.81 / nc + rng.uniform(-1.0, 1.0, (B, 4))
    f[..., 0:2] = rng.uniform(-0.3, 0.3, (B, 4, 2)) * f[..., 2:3]
    f *= contacts[..., None]
    # foot goals around the nominal stance (shoulders, ground at -h_ref)
    goals = np.zeros((B, 3, 4))
    goals[:, 0] = SHOULDERS[0] + rng.uniform(-0.03, 0.03, (B, 4))
    goals[:, 1] = SHOULDERS[1] + rng.uniform(-0.03, 0.03, (B, 4))
    goals[:, 2] = -0.22 + rng.uniform(-0.03, 0.03, (B, 4))
    vgoals = rng.uniform(-0.5, 0.5, (B, 3, 4))
    agoals = rng.uniform(-5.0, 5.0, (B, 3, 4))
    return q, dq, f.reshape(B, 12), contacts, pack_cols(goals), pack_cols(vgoals), pack_cols(agoals)


class PlannerWorkload:
    """Per-tick inputs of the planners for B robots: what scripts/Controller.py:updateState (:383-425) hands to
    Gait / FootstepPlanner / StatePlanner each control tick -- the base pose q.head(7) integrated from the velocity
    reference ("perfect" x, y, yaw; measured height, roll, pitch replaced by seeded noise), the base velocity in the
    horizontal frame h_v (reference + noise) and the joystick reference v_ref, piecewise constant with a few gait
    changes (joystick codes of Gait::changeGait; 5 = walk).  Deterministic in (B, seed); tick(k) must be called for
    k = 0, 1, 2, ... in order."""

    def __init__(self, B, seed, dt_wbc=0.002, k_mpc=10, codes=(1, 2, 3, 4), change_every=170, yaw0=None, speed=1.0):
        self.B, self.dt, self.k_mpc, self.speed = B, dt_wbc, k_mpc, speed
        self.rng = np.random.default_rng(seed)
        self.codes, self.change_every = tuple(codes), change_every
        self.xy = np.zeros((B, 2))
        # yaw0 = None: random initial heading (planner-only tests); closed-loop runs through the whole-body tick start at
        # heading 0, where the feet stand under the shoulders as FootTrajectoryGenerator::initialize assumes
        self.yaw = self.rng.uniform(-3.0, 3.0, B) if yaw0 is None else np.full(B, float(yaw0))
        self.vref = np.zeros((B, 6))
        self._new_vref()
        self.k = 0

    def _new_vref(self):
        B, r = self.B, self.rng
        self.vref[:, 0] = r.uniform(-0.5, 1.5, B)
        self.vref[:, 1] = r.uniform(-0.5, 0.5, B)
        self.vref[:, 5] = r.uniform(-0.7, 0.7, B)
        self.vref[r.random(B) < 0.1, 5] = 0.0  # the yaw-rate == 0 branch of the planners
        self.vref *= self.speed

    def tick(self, k):
        assert k == self.k, "ticks must be consumed in order"
        B, r = self.B, self.rng
        code = np.zeros(B, np.int32)
        if k > 0 and k % self.change_every == 0 and self.codes:
            pick = r.random(B) < 0.5
            code[pick] = r.choice(self.codes, pick.sum())
            self._new_vref()
        # Controller.py:393-403: perfect integration of x, y, yaw
        c, s = np.cos(self.yaw), np.sin(self.yaw)
        self.xy[:, 0] += (c * self.vref[:, 0] - s * self.vref[:, 1]) * self.dt
        self.xy[:, 1] += (s * self.vref[:, 0] + c * self.vref[:, 1]) * self.dt
        self.yaw += self.vref[:, 5] * self.dt
        roll, pitch = r.uniform(-0.1, 0.1, B), r.uniform(-0.1, 0.1, B)
        q7 = np.zeros((B, 7))
        q7[:, 0:2] = self.xy
        q7[:, 2] = r.uniform(0.2, 0.26, B)
        q7[:, 3:7] = euler_to_quaternion(roll, pitch, self.yaw)
        h_v = self.vref + self.speed * r.uniform(-0.1, 0.1, (B, 6))
        self.k += 1
        return q7, h_v, self.vref.copy(), code


def euler_to_quaternion(roll, pitch, yaw):
    """(x, y, z, w) of Rz(yaw) Ry(pitch) Rx(roll) -- Estimator.EulerToQuaternion of the reference scripts."""
    cr, sr, cp, sp, cy, sy = np.cos(roll / 2), np.sin(roll / 2), np.cos(pitch / 2), np.sin(pitch / 2), np.cos(yaw / 2), np.sin(yaw / 2)
    return np.stack([sr * cp * cy - cr * sp * sy, cr * sp * cy + sr * cp * sy, cr * cp * sy - sr * sp * cy,
                     cr * cp * cy + sr * sp * sy], axis=-1)
