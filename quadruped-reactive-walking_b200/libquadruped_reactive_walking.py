"""Drop-in for the reference's Boost.Python module ``libquadruped_reactive_walking``
(/root/reference/python/gepadd.cpp:283-298) on top of the B200 library.

Same class and method names, same argument meaning, numpy matrices in the shapes eigenpy hands to
and from ``Eigen::MatrixXd`` -- so ``scripts/MPC_Wrapper.py:4``, ``scripts/QP_WBC.py:8`` and
``scripts/solo12InvKin.py:4`` can ``import libquadruped_reactive_walking`` unchanged.  Each object
is a batch-of-one view of the batched CUDA solvers (qrw_b200.Batched*); there is no CPU path.

Note for ``MPC_Wrapper``'s multiprocessing mode (scripts/MPC_Wrapper.py:150-225): the MPC object is
created inside the child process, so the CUDA context is created there, after the fork.
"""
import os

import numpy as np

from qrw_b200.batched import BatchedInvKin, BatchedMPC, BatchedQPWBC


def add(a, b):  # gepetto::example::add, src/gepadd.cpp
    return int(a) + int(b)


def sub(a, b):
    return int(a) - int(b)


def _f(a):
    """numpy matrix -> column-major flat doubles (what eigenpy copies into Eigen::MatrixXd)."""
    return np.asarray(a, dtype=np.float64).ravel(order="F")[None, :]


class Params:
    """include/qrw/Params.hpp:37-61 -- the 17 public fields, default values of
    src/config_solo12.yaml; ``initialize(path)`` reads a yaml file with the same keys."""

    def __init__(self):
        self.interface = "enp2s0"
        self.SIMULATION = True
        self.LOGGING = False
        self.PLOTTING = True
        self.dt_wbc = 0.002
        self.N_gait = 20
        self.envID = 0
        self.velID = 2
        self.dt_mpc = 0.02
        self.T_gait = 0.32
        self.T_mpc = 0.32
        self.N_SIMULATION = 3000
        self.type_MPC = True
        self.use_flat_plane = True
        self.predefined_vel = True
        self.kf_enabled = False
        self.enable_pyb_GUI = True

    def initialize(self, file_path):
        import yaml

        if not os.path.exists(file_path):
            raise RuntimeError(f"Error opening the parameters file {file_path}")  # Params.cpp:31-32 assert
        with open(file_path) as fh:
            doc = yaml.safe_load(fh)
        if not isinstance(doc, dict) or "robot" not in doc:
            raise RuntimeError("Error: 'robot' node not found in the yaml file")  # Params.hpp:64-70
        for key in list(vars(self)):
            if key not in doc["robot"]:
                raise RuntimeError(f"Error: parameter '{key}' not found in the yaml file")  # Params.hpp:72-88
            setattr(self, key, type(getattr(self, key))(doc["robot"][key]))


class MPC:
    """class MPC, python/gepadd.cpp:16-43.  MPC(dt, n_steps, T_gait, N_gait) or MPC(params)."""

    def __init__(self, dt_in=None, n_steps_in=None, T_gait_in=None, N_gait=None):
        self._impl = None
        if dt_in is None:
            return  # default constructor builds an empty object, src/MPC.cpp:34
        if isinstance(dt_in, Params):
            p = dt_in
            dt_in, n_steps_in, T_gait_in, N_gait = p.dt_mpc, int(round(p.T_mpc / p.dt_mpc)), p.T_gait, p.N_gait
        self._n, self._N_gait = int(n_steps_in), int(N_gait)
        self._impl = BatchedMPC(1, dt_in, n_steps_in, T_gait_in, N_gait)

    def run(self, num_iter, xref_in, fsteps_in):
        xref_in, fsteps_in = np.asarray(xref_in), np.asarray(fsteps_in)
        if xref_in.shape != (12, self._n + 1) or fsteps_in.shape != (self._N_gait, 12):
            raise ValueError("xref_in must be 12 x (n_steps+1) and fsteps_in N_gait x 12")
        return self._impl.run(int(num_iter), _f(xref_in), _f(fsteps_in))

    def get_latest_result(self):
        return self._impl.get_latest_result().reshape(self._n, 24).T.copy()

    def get_gait(self):
        return self._impl.get_gait().reshape(4, self._N_gait).T.copy()

    def get_Sgait(self):
        return self._impl.get_Sgait().reshape(-1, 1).copy()


class QPWBC:
    """class QPWBC, python/gepadd.cpp:209-238."""

    def __init__(self):
        self._impl = BatchedQPWBC(1)

    def run(self, M, Jc, f_cmd, RNEA, k_contacts=None):
        M, Jc = np.asarray(M), np.asarray(Jc)
        if M.shape != (18, 18) or Jc.shape != (12, 18):
            raise ValueError("M must be 18 x 18 and Jc 12 x 18")
        return self._impl.run(_f(M), _f(Jc), _f(f_cmd), _f(RNEA))

    compute = run

    def get_f_res(self):
        return self._impl.get_f_res().reshape(12, 1).copy()

    def get_ddq_res(self):
        return self._impl.get_ddq_res().reshape(6, 1).copy()

    def get_H(self):
        return self._impl.get_H().reshape(12, 12).T.copy()


class InvKin:
    """class InvKin, python/gepadd.cpp:178-206."""

    def __init__(self, dt_in=0.0):
        self.dt = dt_in
        self._impl = BatchedInvKin(1)

    def refreshAndCompute(self, contacts, goals, vgoals, agoals, posf, vf, wf, af, Jf):
        return self._impl.refreshAndCompute(*[_f(a) for a in (contacts, goals, vgoals, agoals, posf, vf, wf, af, Jf)]).reshape(12, 1).copy()

    def get_q_step(self):
        return self._impl.get_q_step().reshape(12, 1).copy()

    def get_dq_cmd(self):
        return self._impl.get_dq_cmd().reshape(12, 1).copy()


# ------------------------------------------------------------------------------------------------------------------
# Planner classes (python/gepadd.cpp:45-175).  The reference wires them by reference: FootstepPlanner and
# FootTrajectoryGenerator hold a pointer to the Gait object.  Here the Gait object owns a batch-of-one planner handle
# (qrw_b200.planner.BatchedPlanner) and the other classes attach to it; every method runs the matching stage of the one
# CUDA kernel.  A StatePlanner built on its own gets its own handle.
# ------------------------------------------------------------------------------------------------------------------
def _m34(a):
    """[1,12] foot-major (3 x 4 column-major) -> 3 x 4 matrix."""
    return np.asarray(a, dtype=np.float64).reshape(4, 3).T.copy()


class _PlannerCore:
    """Shared batch-of-one planner handle with the last inputs each stage was called with."""

    def __init__(self, dt_mpc, dt_wbc, T_gait, T_mpc, N_gait, k_mpc, h_ref):
        from qrw_b200.planner import BatchedPlanner

        self.args = dict(dt_mpc=dt_mpc, dt_wbc=dt_wbc, T_gait=T_gait, T_mpc=T_mpc, N_gait=N_gait, k_mpc=k_mpc, h_ref=h_ref)
        self.impl = BatchedPlanner(1, **self.args)
        self.cfg = dict(shoulders=None, ftg_target=None, ftg_position=None, max_height=0.05, lock_time=0.07)
        self.q7 = np.array([[0.0, 0.0, h_ref, 0.0, 0.0, 0.0, 1.0]])
        self.v6 = np.zeros((1, 6))
        self.vref6 = np.zeros((1, 6))

    def rebuild(self, **kw):
        """initialize() of a later class changed a constructor-time constant: new handle, same configuration."""
        from qrw_b200.planner import BatchedPlanner

        self.args.update(kw)
        self.impl.close()
        self.impl = BatchedPlanner(1, **self.args)
        self.impl.configure(**self.cfg)

    def configure(self, **kw):
        self.cfg.update(kw)
        self.impl.configure(**self.cfg)


class Gait:
    """class Gait, python/gepadd.cpp:76-109 (src/Gait.cpp)."""

    def __init__(self):
        self._core = None

    def initialize(self, dt_in, T_gait_in, T_mpc_in, N_gait):
        from qrw_b200 import QrwError

        try:
            # dt_wbc / k_mpc / h_ref are set by the classes that own them (FootstepPlanner / FootTrajectoryGenerator)
            self._core = _PlannerCore(dt_in, dt_in / 10.0, T_gait_in, T_mpc_in, int(N_gait), 10, 0.0)
        except QrwError as e:  # Gait.cpp:30-31 throws std::invalid_argument, which Boost.Python maps to ValueError
            raise ValueError(str(e)) from None
        self._N = int(N_gait)

    def updateGait(self, k, k_mpc, q, joystickCode):
        c = self._core
        if int(k_mpc) != c.args["k_mpc"]:
            raise ValueError("k_mpc differs from the one the planners were initialised with")
        c.q7 = np.asarray(q, dtype=np.float64).reshape(-1)[:7][None, :].copy()
        c.impl.run(int(k), c.q7, c.v6, c.vref6, [int(joystickCode)], stages=1)

    def getCurrentGait(self):
        return self._core.impl.outputs(("gait",))["gait"].reshape(4, self._N).T.copy()

    def getPastGait(self):
        return self._core.impl.state()["past"].reshape(4, self._N).T.copy()

    def getDesiredGait(self):
        return self._core.impl.state()["desired"].reshape(4, self._N).T.copy()

    def isNewPhase(self):
        return bool(self._core.impl.state()["flags"][0, 0])

    def getIsStatic(self):
        return bool(self._core.impl.state()["flags"][0, 1])

    def setGait(self, gaitMatrix):  # Gait.cpp:262-269: prints the matrix and returns false
        print("Gait matrix received by setGait:")
        print(np.asarray(gaitMatrix))
        return False


class StatePlanner:
    """class StatePlanner, python/gepadd.cpp:45-73 (src/StatePlanner.cpp)."""

    def __init__(self):
        self._impl = None

    def initialize(self, dt_in, T_mpc_in, h_ref_in):
        from qrw_b200.planner import BatchedPlanner

        self._n = int(round(T_mpc_in / dt_in))
        N = max(self._n + 4, 20)
        self._impl = BatchedPlanner(1, dt_mpc=dt_in, dt_wbc=dt_in / 10.0, T_gait=T_mpc_in, T_mpc=T_mpc_in, N_gait=N + N % 2, k_mpc=10,
                                    h_ref=h_ref_in)
        self._z_warned = False

    def computeReferenceStates(self, q, v, b_vref, z_average):
        if float(z_average) != 0.0:
            raise ValueError("z_average != 0 is not supported (the reference always passes 0.0, Controller.py:235-236)")
        q7 = np.asarray(q, dtype=np.float64).reshape(-1)[:7][None, :]
        self._impl.run(0, q7, np.asarray(v, np.float64).reshape(1, 6), np.asarray(b_vref, np.float64).reshape(1, 6), None, stages=8)

    def getReferenceStates(self):
        return self._impl.outputs(("xref",))["xref"].reshape(self._n + 1, 12).T.copy()

    def getNSteps(self):
        return self._n


class FootstepPlanner:
    """class FootstepPlanner, python/gepadd.cpp:112-141 (src/FootstepPlanner.cpp)."""

    def __init__(self):
        self._core = None

    def initialize(self, dt_in, dt_wbc_in, T_mpc_in, h_ref_in, shouldersIn, gaitIn, N_gait):
        c = self._core = gaitIn._core
        self._N = int(N_gait)
        sh = np.asarray(shouldersIn, dtype=np.float64)[:3, :4]
        c.cfg.update(shoulders=sh)
        c.rebuild(dt_wbc=float(dt_wbc_in), h_ref=float(h_ref_in), k_mpc=int(round(dt_in / dt_wbc_in)))

    def updateFootsteps(self, refresh, k, q, b_v, b_vref):
        """`k` is the remaining number of control ticks before the next MPC tick and `refresh` the new-MPC-tick flag, as
        Controller.py:225-227 computes them from the tick index; the kernel takes the tick index itself."""
        c = self._core
        k_mpc = c.args["k_mpc"]
        tick = (k_mpc - int(k)) % k_mpc  # k = k_mpc - tick % k_mpc
        if bool(refresh) and tick == 0:
            tick = k_mpc  # any non-zero multiple of k_mpc: refresh = (tick % k_mpc == 0 and tick != 0)
        elif bool(refresh):
            raise ValueError("refresh is only meaningful on an MPC tick (k == k_mpc)")
        c.q7 = np.asarray(q, dtype=np.float64).reshape(-1)[:7][None, :].copy()
        c.v6 = np.asarray(b_v, dtype=np.float64).reshape(1, 6).copy()
        c.vref6 = np.asarray(b_vref, dtype=np.float64).reshape(1, 6).copy()
        c.impl.run(tick, c.q7, c.v6, c.vref6, None, stages=2)
        return _m34(c.impl.outputs(("o_target",))["o_target"])

    def getFootsteps(self):
        return self._core.impl.outputs(("fsteps",))["fsteps"].reshape(12, self._N).T.copy()

    def getTargetFootsteps(self):
        return _m34(self._core.impl.outputs(("o_target",))["o_target"])

    def getRz(self):
        """Rz as the last statement that wrote it leaves it (FootstepPlanner.cpp:213-214: rotation by the base yaw)."""
        x, y, z, w = self._core.q7[0, 3:7]
        yaw = np.arctan2(2 * (x * y + z * w), 1 - 2 * (y * y + z * z))
        c, s = np.cos(yaw), np.sin(yaw)
        return np.array([[c, -s, 0.0], [s, c, 0.0], [0.0, 0.0, 1.0]])


class FootTrajectoryGenerator:
    """class FootTrajectoryGenerator, python/gepadd.cpp:144-175 (src/FootTrajectoryGenerator.cpp)."""

    def __init__(self):
        self._core = None

    def initialize(self, maxHeightIn, lockTimeIn, targetFootstepIn, initialFootPosition, dt_tsid_in, k_mpc_in, gaitIn):
        c = self._core = gaitIn._core
        c.cfg.update(ftg_target=np.asarray(targetFootstepIn, np.float64)[:3, :4], ftg_position=np.asarray(initialFootPosition, np.float64)[:3, :4],
                     max_height=float(maxHeightIn), lock_time=float(lockTimeIn))
        c.rebuild(dt_wbc=float(dt_tsid_in), k_mpc=int(k_mpc_in))

    def update(self, k, targetFootstep):
        c = self._core
        t = np.asarray(targetFootstep, dtype=np.float64)[:3, :4]
        c.impl.set_target(np.ascontiguousarray(t.T.reshape(1, 12)))
        c.impl.run(int(k), c.q7, c.v6, c.vref6, None, stages=4)

    def _get(self, name):
        return _m34(self._core.impl.outputs((name,))[name])

    def getFootPosition(self):
        return self._get("fpos")

    def getFootVelocity(self):
        return self._get("fvel")

    def getFootAcceleration(self):
        return self._get("facc")
