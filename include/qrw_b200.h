/* qrw_b200.h -- C ABI of the B200-native batched controller core (libqrw_b200.so).
 *
 * Drop-in boundary for the per-tick hot path of paLeziart/quadruped-reactive-walking.  Each
 * entry point replaces a method the reference exposes through Boost.Python
 * (/root/reference/python/gepadd.cpp) on the C++ classes MPC, QPWBC and InvKin; the reference
 * handles ONE robot per object on the CPU, these handle a batch of B independent robots per
 * object on one GPU.  INTEGRATION.md shows the binding a maintainer would add on the
 * reference side.
 *
 * Conventions
 *  - plain C: opaque handles, pointers and sizes, int return codes (0 = ok, non-zero = error,
 *    text from qrw_last_error()).  No C++ or torch types, no exceptions across the boundary.
 *  - every array is instance-major; inside one instance matrices are COLUMN-MAJOR doubles,
 *    i.e. exactly what Eigen::MatrixXd::data() holds in the reference:
 *        xref   [B][12*(n_steps+1)]   element (r,c) at r + 12*c          (MPC.cpp:626 xref_in)
 *        fsteps [B][N_gait*12]        element (k,j) at k + N_gait*j      (MPC.cpp:626 fsteps_in)
 *        result [B][24*n_steps]       element (r,i) at r + 24*i          (MPC.cpp:604)
 *  - "_device" entry points take device pointers and are asynchronous on the given CUDA stream
 *    (pass a cudaStream_t cast to void*, NULL = default stream); "_host" entry points take host
 *    pointers, do the H2D / D2H copies themselves and return when the data is there.
 *  - there is no CPU fallback: every call fails with an error if no CUDA device is usable.
 */
#ifndef QRW_B200_H
#define QRW_B200_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct qrw_mpc qrw_mpc_t;
typedef struct qrw_wbc qrw_wbc_t;
typedef struct qrw_wbctick qrw_wbctick_t;

const char* qrw_last_error(void);
/* number of kernels this library has launched since load (bench.py's gpu_launches) */
long long qrw_launch_count(void);
int qrw_device_sm_count(int device, int* sm_count);
/* measured FP64 FMA throughput of `device` in TFLOP/s (DFMA microbenchmark; the denominator of the
 * FP64 roofline fraction bench.py reports next to the HBM one) */
int qrw_fp64_peak(int device, double* tflops);

/* ---------------------------------------------------------------------------------------
 * MPC -- replaces class MPC (include/qrw/MPC.hpp:16-129, src/MPC.cpp)
 * ------------------------------------------------------------------------------------- */
/* MPC::MPC(dt, n_steps, T_gait, N_gait), src/MPC.cpp:3-32, for `batch` robots on `device`.
 * LIMIT: only the reference configuration n_steps == 16 (T_mpc / dt_mpc = 0.32 / 0.02, src/config_solo12.yaml:11-13)
 * is compiled in -- the kernels map one horizon stage to 16 lanes of a 256-thread CTA; any other value fails with a
 * message naming the limit.  n_steps < N_gait <= 32. */
int qrw_mpc_create(qrw_mpc_t** out, int batch, double dt, int n_steps, double T_gait, int N_gait, int device);
int qrw_mpc_destroy(qrw_mpc_t* h);
/* MPC::run(num_iter, xref_in, fsteps_in), src/MPC.cpp:626-649.  num_iter == 0 takes the
 * create_matrices/osqp_setup path (cold start), anything else the update path (warm start from
 * the state left by the previous call).  One value for the whole batch.
 * STREAM CONTRACT: _device entry points are asynchronous on the caller's `stream`.  Every launch is followed by an
 * event on that stream, and qrw_mpc_result_host / qrw_mpc_gait_host / qrw_mpc_sgait_host wait for the event of the
 * LAST launch before they copy (they block the host until the copy is done), so run_device(...) followed by
 * result_host(...) needs no synchronisation by the caller.  qrw_mpc_info_host and the debug / wbc / wbctick host
 * getters synchronise the whole device.  Exception: a launch captured into a CUDA graph records no event -- callers
 * that replay graphs synchronise the replay stream themselves.  qrw_mpc_result_device only returns a pointer: reading
 * it is ordered by the caller's own stream. */
int qrw_mpc_run_device(qrw_mpc_t* h, int num_iter, const double* d_xref, const double* d_fsteps, void* stream);
/* Host entry point (synchronous).  Pageable inputs are staged with two host-to-device copies.  PINNED inputs (cudaMallocHost,
 * cudaHostRegister, torch pin_memory, the handle's own buffers of qrw_mpc_pinned) take the ZERO-COPY path: the kernel reads a
 * robot's xref / fsteps straight from host memory (unified addressing) when it starts on that robot and mirrors the robot's
 * result into the handle's pinned result buffer when it is done, so both transfers run under the solves of the other robots
 * instead of in front of and behind the launch (config 2, end to end: 158.8 k against 155.8 k solves/s).
 * qrw_mpc_result_host(h, <that pinned result buffer>) then has nothing left to copy.  The inputs must stay unchanged until
 * the call returns (it does not return before the solves are done). */
int qrw_mpc_run_host(qrw_mpc_t* h, int num_iter, const double* xref, const double* fsteps);
/* MPC::run with a per-robot call index d_k [B] (device ints): < 0 = robot not called in this launch (solver state and
 * result untouched), 0 = cold start, > 0 = update path.  For servers whose robots post requests at different times. */
int qrw_mpc_run_indexed_device(qrw_mpc_t* h, const int* d_k, const double* d_xref, const double* d_fsteps, void* stream);
/* MPC::get_latest_result(), src/MPC.cpp:604: [B][24*n_steps] */
int qrw_mpc_result_host(qrw_mpc_t* h, double* out);
int qrw_mpc_result_device(qrw_mpc_t* h, const double** d_out);
/* MPC::get_gait() / get_Sgait(), src/MPC.cpp:770-780: [B][N_gait*4] and [B][12*n_steps] doubles */
int qrw_mpc_gait_host(qrw_mpc_t* h, double* out);
int qrw_mpc_sgait_host(qrw_mpc_t* h, double* out);
/* what OSQP keeps in work->info and the reference throws away (src/MPC.cpp:558-563):
 * per-instance ADMM iterations, OSQP status value, rho after the solve, KKT factorisations.
 * Any pointer may be NULL. */
int qrw_mpc_info_host(qrw_mpc_t* h, int* iters, int* status, double* rho, int* n_factor);
/* test hooks: iteration cap (default 4000, OSQP max_iter) and a dump of the Ruiz scaling
 * (D[384] | E[704] | c | unscaled dynamics bounds[192]) of every instance after the next run */
int qrw_mpc_set_max_iter(qrw_mpc_t* h, int max_iter);
int qrw_mpc_debug_host(qrw_mpc_t* h, double* out /* [B][1281] */);
/* which CUDA kernel runs the solve.  The shipped library contains ONE: 2 = force-first elimination, one 256-thread CTA per
 * instance, matrix rows in tensor memory (mpc_v2_core.cuh); any other value fails.  The A/B test build of the same source
 * (-DQRW_AB_KERNELS, tests/ab) also accepts 4 = state chain over super-stages, 3 = nested-dissection state solve, 1 = 24x24
 * block-tridiagonal CTA kernel, 0 = one warp per instance (superseded / not adopted; same persistent state, same OSQP iterate
 * sequence). */
int qrw_mpc_set_kernel(qrw_mpc_t* h, int variant);
/* work order of the persistent grid: 1 (default) = robots sorted by the expected length of their solve, longest first --
 * max(ADMM iteration count of the previous solve, 2000 x commanded horizontal speed xref[6:8, 1] in m/s) -- so that the long
 * solves start early and the grid drains evenly (measured on B200, config 2: 158.6 k against 149.3 k solves/s); 0 = index
 * order.  Results do not depend on it (each robot's solve is unchanged, only when it starts). */
int qrw_mpc_set_work_order(qrw_mpc_t* h, int longest_first);
/* The solver state carried from call to call (scaled ADMM iterates, rho, stale B_k values, gait,
 * S_gait -- the reference keeps it inside OSQP's workspace and the MPC object).  Snapshot and
 * restore, e.g. to replay a benchmark from the same history or to checkpoint a fleet. */
int qrw_mpc_state_bytes(qrw_mpc_t* h, size_t* bytes);
int qrw_mpc_state_save(qrw_mpc_t* h, void* d_buf, void* stream);
int qrw_mpc_state_load(qrw_mpc_t* h, const void* d_buf, void* stream);
/* pinned host buffers owned by the handle (inputs [B][12 * 17], [B][12 * N_gait]; result [B][24 * 16]): what the zero-copy
 * path of qrw_mpc_run_host / qrw_mpc_result_host reads and writes */
int qrw_mpc_pinned(qrw_mpc_t* h, double** xref, double** fsteps, double** result);

/* ---------------------------------------------------------------------------------------
 * QPWBC -- replaces class QPWBC (include/qrw/QPWBC.hpp:16-98, src/QPWBC.cpp)
 * ------------------------------------------------------------------------------------- */
int qrw_wbc_create(qrw_wbc_t** out, int batch, int device);
int qrw_wbc_destroy(qrw_wbc_t* h);
/* QPWBC::run(M, Jc, f_cmd, RNEA, k_contact), src/QPWBC.cpp:310-343 (k_contact is accepted and
 * ignored by the reference, :345-362; it is not part of this ABI).
 * M [B][18*18], Jc [B][12*18], f_cmd [B][12], RNEA [B][6]. */
int qrw_wbc_run_device(qrw_wbc_t* h, const double* d_M, const double* d_Jc, const double* d_f_cmd,
                       const double* d_RNEA, void* stream);
int qrw_wbc_run_host(qrw_wbc_t* h, const double* M, const double* Jc, const double* f_cmd, const double* RNEA);
/* get_f_res [B][12], get_ddq_res [B][6], get_H [B][144] (src/QPWBC.cpp:302-308); NULL skips */
int qrw_wbc_result_host(qrw_wbc_t* h, double* f_res, double* ddq_res, double* H);
int qrw_wbc_result_device(qrw_wbc_t* h, const double** d_f_res, const double** d_ddq_res);
int qrw_wbc_info_host(qrw_wbc_t* h, int* iters, int* status, double* rho);

/* ---------------------------------------------------------------------------------------
 * InvKin -- replaces InvKin::refreshAndCompute (src/InvKin.cpp:23-73); stateless.
 * contacts [B][4]; goals, vgoals, agoals [B][3*4]; posf, vf, wf, af [B][4*3]; Jf [B][12*12];
 * outputs ddq (return value), dq_cmd (get_dq_cmd), q_step (get_q_step): [B][12] each.
 * ------------------------------------------------------------------------------------- */
int qrw_invkin_run_device(int batch, const double* contacts, const double* goals, const double* vgoals,
                          const double* agoals, const double* posf, const double* vf, const double* wf,
                          const double* af, const double* Jf, double* ddq, double* dq_cmd, double* q_step,
                          void* stream);
int qrw_invkin_run_host(int batch, const double* contacts, const double* goals, const double* vgoals,
                        const double* agoals, const double* posf, const double* vf, const double* wf,
                        const double* af, const double* Jf, double* ddq, double* dq_cmd, double* q_step);

/* ---------------------------------------------------------------------------------------
 * Whole-body tick -- replaces wbc_controller.compute (scripts/QP_WBC.py:52-131) for B robots: the Pinocchio
 * calls around the C++ classes (fixed-base foot kinematics solo12InvKin.py:47-59, contact Jacobians
 * QP_WBC.py:96-101, rnea :104,:116, constant crba diagonal :89-93), InvKin::refreshAndCompute, QPWBC::run and
 * tau_ff = rnea(q, dq, ddq_cmd + [ddq_res; 0])[6:] - Jc[:, 6:]' f_res (:110-117).
 * q [B][19] (xyz, quaternion xyzw, 12 joints), dq [B][18] (base twist in the base frame, 12 joints),
 * f_cmd [B][12], contacts [B][4], pgoals / vgoals / agoals [B][12] (3 x 4 column-major);
 * outputs [B][12]: tau_ff, qdes (joints: q + q_step), vdes (dq_cmd), f_res (f_with_delta), ddq_cmd (joints).
 * Solo12 constants: include/solo12_model.h (not in the reference tree: PARITY UNPINNED at this sub-boundary).
 * ------------------------------------------------------------------------------------- */
int qrw_wbctick_create(qrw_wbctick_t** out, int batch, int device);
int qrw_wbctick_destroy(qrw_wbctick_t* h);
int qrw_wbctick_run_device(qrw_wbctick_t* h, const double* q, const double* dq, const double* f_cmd,
                           const double* contacts, const double* pgoals, const double* vgoals, const double* agoals,
                           void* stream);
/* same with f_cmd read with a stride (f_cmd + b * f_stride, f_stride >= 12): the MPC result + 12 with stride 24 * n_steps
 * feeds x_f_mpc[12:24, 0] (Controller.py:258) without a gather */
int qrw_wbctick_run_fstride_device(qrw_wbctick_t* h, const double* q, const double* dq, const double* f_cmd, int f_stride,
                                   const double* contacts, const double* pgoals, const double* vgoals, const double* agoals,
                                   void* stream);
int qrw_wbctick_run_host(qrw_wbctick_t* h, const double* q, const double* dq, const double* f_cmd, const double* contacts,
                         const double* pgoals, const double* vgoals, const double* agoals);
int qrw_wbctick_result_host(qrw_wbctick_t* h, double* tau_ff, double* qdes, double* vdes, double* f_res, double* ddq_cmd);
int qrw_wbctick_result_device(qrw_wbctick_t* h, const double** tau_ff, const double** qdes, const double** vdes,
                              const double** f_res);
/* The tick as the control loop calls it (scripts/Controller.py:275-303): q_wbc = (0, 0, h_ref, 0, 0, 0, 1, previous qdes),
 * b_v = (v_ref[0:6], previous vdes) are built on the device from the handle's own previous outputs; f_cmd is read with a
 * stride (the MPC result + 12, stride 24 * n_steps = x_f_mpc[12:24, 0], Controller.py:258,300).  set_desired initialises
 * qdes / vdes (Controller.py:153; host pointers [B][12], NULL = zeros). */
int qrw_wbctick_set_desired_host(qrw_wbctick_t* h, const double* qdes, const double* vdes);
int qrw_wbctick_run_ctrl_device(qrw_wbctick_t* h, double h_ref, const double* vref6, const double* f_cmd, int f_stride,
                                const double* contacts, const double* pgoals, const double* vgoals, const double* agoals,
                                void* stream);
/* intermediates handed to QPWBC (Jc [B][216], RNEA [B][6]) and its ddq_res [B][6]; any pointer may be NULL.  Jc / RNEA stay
 * on the SM unless qrw_wbctick_set_debug(h, 1) was called before the tick (they are then copied to HBM as well). */
int qrw_wbctick_set_debug(qrw_wbctick_t* h, int on);
/* robots per warp of the tick kernel (csrc/wbc_tick_core.cuh): 2, 4 or 8; 0 (default) = chosen at launch from the batch
 * size.  A tuning knob only: results do not depend on it. */
int qrw_wbctick_set_group(qrw_wbctick_t* h, int group);
int qrw_wbctick_debug_host(qrw_wbctick_t* h, double* Jc, double* rnea6, double* ddq_res);
/* ONE whole-body tick that allocates NOTHING.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same kernel and same arguments as qrw_wbctick_run_device, but without a handle: the kernel's scratch (QP state, ddq_res,
 * ddq_cmd, info, work counter, pinv(M)) lives in a file-scope __device__ array of fixed size (csrc/noalloc_abi.cuh, 3.0 MB,
 * held by every GPU that loads the library until the process exits); the call issues one cudaMemsetAsync of 4 bytes and one
 * kernel launch on `stream` and calls no cudaMalloc / cudaFree / allocating routine.  All pointers are DEVICE pointers of the
 * caller: q [B][19], dq [B][18], f_cmd [B][12], contacts [B][4], goals [B][12] in; tau_ff, qdes, vdes, f_res [B][12] out.
 * LIMITS: batch <= QRW_WBCTICK_NOALLOC_MAX_ROBOTS, anything larger is rejected with an error; the QP is always cold-started
 * (a call equals the FIRST qrw_wbctick_run_device of a fresh handle -- a warm start needs state between calls, i.e. a
 * handle); every call on a device shares the one array, which keeps the last call's contents (the call clears what must start
 * at zero): calls that OVERLAP on the GPU race, serialise them; the first call on a device does a one-time set-up
 * (cudaGetSymbolAddress, constants, one device synchronisation): make it outside graph capture. */
#define QRW_WBCTICK_NOALLOC_MAX_ROBOTS 4096
int qrw_wbctick_noalloc_device(int batch, int device, const double* q, const double* dq, const double* f_cmd, const double* contacts,
                               const double* pgoals, const double* vgoals, const double* agoals, double* tau_ff, double* qdes,
                               double* vdes, double* f_res, void* stream);
/* diag(crba(q = 0)[:6, :6]): (m, m, m, Ixx, Iyy, Izz) of the whole robot about the base frame */
int qrw_solo12_base_inertia(double* Y6);
/* what solo12InvKin.py:47-59 reads from Pinocchio: q12, dq12 [B][12] -> posf, vf, wf, af [B][4*3], Jf [B][12*12] */
int qrw_solo12_feet_host(int batch, const double* q12, const double* dq12, double* posf, double* vf, double* wf,
                         double* af, double* Jf);

/* ---------------------------------------------------------------------------------------
 * Planners -- the input producers of the MPC and of the whole-body tick for B robots:
 *   Gait::updateGait                       src/Gait.cpp:189-260           (stage bit 1)
 *   FootstepPlanner::updateFootsteps       src/FootstepPlanner.cpp:52-75  (stage bit 2)
 *   FootTrajectoryGenerator::update        src/FootTrajectoryGenerator.cpp:110-151 (stage bit 4)
 *   StatePlanner::computeReferenceStates   src/StatePlanner.cpp:22-60     (stage bit 8)
 *   feet commands in the base frame        scripts/Controller.py:286-296  (stage bit 16)
 * One call = one control tick k of scripts/Controller.py:222-240 with the stages selected by the mask, in that order;
 * xref / fsteps come out in the layout qrw_mpc_run_device reads, contacts / p_cmd / v_cmd / a_cmd in the layout
 * qrw_wbctick_run_device reads, so a closed-loop batch never leaves HBM.
 * ------------------------------------------------------------------------------------- */
typedef struct qrw_planner qrw_planner_t;
#define QRW_PLAN_GAIT 1
#define QRW_PLAN_FOOTSTEPS 2
#define QRW_PLAN_FOOTTRAJ 4
#define QRW_PLAN_STATE 8
#define QRW_PLAN_FEETCMD 16
#define QRW_PLAN_ALL 31
/* Gait::initialize(dt_mpc, T_gait, T_mpc, N_gait) (Gait.cpp:21-36: trot, fails where the reference throws),
 * StatePlanner::initialize(dt_mpc, T_mpc, h_ref), FootstepPlanner::initialize(dt_mpc, dt_wbc, T_mpc, h_ref, shoulders, gait,
 * N_gait), FootTrajectoryGenerator::initialize(0.05, 0.07, shoulders, shoulders, dt_wbc, k_mpc, gait) with the shoulders
 * of scripts/Controller.py:131-139. */
int qrw_planner_create(qrw_planner_t** out, int batch, double dt_mpc, double dt_wbc, double T_gait, double T_mpc, int N_gait,
                       int k_mpc, double h_ref, int device);
int qrw_planner_destroy(qrw_planner_t* h);
/* other initialize() arguments (host pointers to 3 x 4 column-major matrices shared by the batch, NULL keeps the current
 * value); resets every robot to its initial state */
int qrw_planner_configure(qrw_planner_t* h, const double* shoulders, const double* ftg_target, const double* ftg_position,
                          double max_height, double lock_time);
/* q7 [B][7] = q.head(7) (x y z, quaternion x y z w), v6 [B][6] = h_v, vref6 [B][6], code [B] joystick codes of
 * Gait::changeGait (NULL = 0; 5 = Gait::create_walk, an extension). */
int qrw_planner_run_device(qrw_planner_t* h, int k, int stages, const double* q7, const double* v6, const double* vref6,
                           const int* code, void* stream);
int qrw_planner_run_host(qrw_planner_t* h, int k, int stages, const double* q7, const double* v6, const double* vref6,
                         const int* code);
/* StatePlanner::getReferenceStates [B][12*(n+1)], FootstepPlanner::getFootsteps [B][N_gait*12], currentGait_.row(0)
 * [B][4], o_targetFootstep_, FootTrajectoryGenerator::getFootPosition / Velocity / Acceleration, feet_p/v/a_cmd: [B][12]
 * (3 x 4 column-major).  Any pointer may be NULL. */
int qrw_planner_outputs_device(qrw_planner_t* h, const double** xref, const double** fsteps, const double** contacts,
                               const double** o_target, const double** fpos, const double** fvel, const double** facc,
                               const double** p_cmd, const double** v_cmd, const double** a_cmd);
/* same on the host; gait = Gait::getCurrentGait [B][N_gait*4] */
int qrw_planner_outputs_host(qrw_planner_t* h, double* xref, double* fsteps, double* gait, double* o_target, double* fpos,
                             double* fvel, double* facc, double* p_cmd, double* v_cmd, double* a_cmd);
/* internal state: pastGait_ / desiredGait_ [B][N_gait*4], currentFootstep_ [B][12], t0s / t_swing [B][4],
 * flags [B][2] = (Gait::isNewPhase, Gait::getIsStatic) */
int qrw_planner_state_host(qrw_planner_t* h, double* past_gait, double* desired_gait, double* current_footstep, double* t0s,
                           double* t_swing, int* flags);
/* the `targetFootstep` argument of FootTrajectoryGenerator::update for callers that run stage 4 on its own: [B][12] */
int qrw_planner_set_target_host(qrw_planner_t* h, const double* o_target);

/* ---------------------------------------------------------------------------------------
 * Closed control tick -- scripts/Controller.py:222-303 (Controller.compute) for B robots, entirely in HBM:
 * planners -> MPC::run every k_mpc-th tick (MPC_Wrapper synchronous mode, scripts/MPC_Wrapper.py:64-71,122-148: the very
 * first tick uses the default forces [0, 0, 8] per foot) -> whole-body tick with f_cmd = x_f_mpc[12:24, 0]
 * (Controller.py:258,300), contacts = gait row 0 and the feet commands of :286-296.  What crosses PCIe per robot and
 * tick: 156 B in (q.head(7), h_v, v_ref, joystick code), 288 B out (tau_ff, q_des, v_des).
 * k_mpc = dt_mpc / dt_wbc, n_steps = T_mpc / dt_mpc (16 only, see qrw_mpc_create); q_init12 = wbc_controller.qdes[7:]
 * before the first tick (Controller.py:153; host pointer to 12 doubles shared by the batch, NULL = the Solo12 stance
 * of scripts/main_solo12_control.py).
 * The tick index enters the kernels only through k % k_mpc and k != 0: qrw_tick_capture_period records the k_mpc ticks of
 * one MPC period into a CUDA graph once, qrw_tick_replay_period / qrw_tick_period_host run one period per graph launch
 * from per-period staging buffers (inputs [k_mpc][B][7 | 6 | 6], codes [k_mpc][B], outputs [k_mpc][B][12]).
 * ------------------------------------------------------------------------------------- */
typedef struct qrw_tick qrw_tick_t;
int qrw_tick_create(qrw_tick_t** out, int batch, double dt_wbc, double dt_mpc, double T_gait, double T_mpc, int N_gait, double h_ref,
                    const double* q_init12, int device);
int qrw_tick_destroy(qrw_tick_t* h);
/* one control tick k: q7 [B][7] (x y z, quaternion x y z w), h_v [B][6], v_ref [B][6], code [B] (NULL = 0) */
int qrw_tick_run_device(qrw_tick_t* h, int k, const double* q7, const double* h_v, const double* v_ref, const int* code, void* stream);
int qrw_tick_run_host(qrw_tick_t* h, int k, const double* q7, const double* h_v, const double* v_ref, const int* code);
/* results of the latest tick, [B][12] each (any pointer may be NULL) */
int qrw_tick_result_host(qrw_tick_t* h, double* tau_ff, double* q_des, double* v_des, double* f_res);
int qrw_tick_result_device(qrw_tick_t* h, const double** tau_ff, const double** q_des, const double** v_des, const double** f_res);
/* the planner / MPC / whole-body-tick objects the tick is composed of (getters, state snapshots); owned by the tick handle */
int qrw_tick_parts(qrw_tick_t* h, qrw_planner_t** planner, qrw_mpc_t** mpc, qrw_wbctick_t** wbc);
int qrw_tick_capture_period(qrw_tick_t* h);
int qrw_tick_period_buffers(qrw_tick_t* h, double** q7, double** h_v, double** v_ref, int** code, const double** tau_ff,
                            const double** q_des, const double** v_des);
int qrw_tick_replay_period(qrw_tick_t* h, void* stream);
/* one MPC period from host memory: H2D of [k_mpc][B][.] inputs, one graph launch, D2H of [k_mpc][B][12] outputs */
int qrw_tick_period_host(qrw_tick_t* h, const double* q7, const double* h_v, const double* v_ref, const int* code, double* tau_ff,
                         double* q_des, double* v_des);

/* ---------------------------------------------------------------------------------------
 * MPC server -- the asynchronous mode of scripts/MPC_Wrapper.py (:47-52 shared Values / Arrays, :150-225 the child's
 * loop, :227-264 the wire format) for B robot processes feeding one GPU.  Slot b of the shared block:
 *   data_in  [1 + 12 (n+1) + 12 N_gait] doubles = k / k_mpc | xref C order | fsteps C order   (compress_dataIn, :227-243)
 *   data_out [24 n] doubles, F order                                                            (convert_dataOut, :260-264)
 *   new_data / new_result: signed chars ('b' Values)
 * poll = one pass of the child's loop body for every slot that has new data (claimed slots are solved in ONE launch of
 * the indexed MPC kernel); loop = `while running.value`.  All pointers are host pointers, normally into one shared-memory
 * block mapped by the robot processes (qrw_b200/mpc_server.py).
 * ------------------------------------------------------------------------------------- */
typedef struct qrw_mpc_server qrw_mpc_server_t;
int qrw_mpc_server_create(qrw_mpc_server_t** out, qrw_mpc_t* mpc);
int qrw_mpc_server_destroy(qrw_mpc_server_t* s);
int qrw_mpc_server_slot_doubles(qrw_mpc_server_t* s, int* per_in, int* per_out);
int qrw_mpc_server_poll_host(qrw_mpc_server_t* s, const double* data_in, signed char* new_data, double* data_out,
                             signed char* new_result, int* n_served);
int qrw_mpc_server_loop_host(qrw_mpc_server_t* s, const double* data_in, signed char* new_data, double* data_out,
                             signed char* new_result, const signed char* running, int idle_us, long long max_rounds,
                             long long* total_served);

#ifdef __cplusplus
}
#endif
#endif /* QRW_B200_H */
