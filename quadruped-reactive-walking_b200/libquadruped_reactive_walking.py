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
