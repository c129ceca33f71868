"""TEST INFRASTRUCTURE: the product's MPC kernel core (csrc/mpc_v2_core.cuh) compiled by g++ in CTA emulation (tests/emu),
behind the host interface of qrw_b200.BatchedMPC, so that host code written against BatchedMPC -- the drop-in module, the
reference's own scripts on top of it -- can be exercised on a machine without a GPU.  Never shipped, never a fallback."""
import ctypes as C

import numpy as np


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class EmuMPC:
    def __init__(self, emu, B, N_gait=20, core="v2"):
        self.L, self.B, self.N_gait = emu, B, N_gait
        # "cta": csrc/mpc_cta_core.cuh (256-thread CTA per instance); "warp": csrc/mpc_core.cuh
        # "v2": csrc/mpc_v2_core.cuh (force-first elimination, twisted state chain: the default kernel)
        # "v3": csrc/mpc_v3_core.cuh (force-first elimination, nested-dissection state solve: A/B variant)
        self.fn = {"v4": emu.emu_mpc_v4_run, "v3": emu.emu_mpc_v3_run, "v2": emu.emu_mpc_v2_run, "cta": emu.emu_mpc_cta_run, "warp": emu.emu_mpc_run}[core]
        self.std = np.zeros((B, emu.emu_mpc_state_doubles()))
        self.sti = np.zeros((B, emu.emu_mpc_state_ints()), np.int32)
        self.out = np.zeros((B, 384))
        self.info = np.zeros((B, 4), np.int32)
        self.rho = np.zeros(B)
        self.dbg = np.zeros((B, 1281))

    def run(self, k, xref, fsteps, threads=8, max_iter=0):
        xref, fsteps = np.ascontiguousarray(xref, dtype=np.float64), np.ascontiguousarray(fsteps, dtype=np.float64)
        self.fn(self.B, C.c_double(0.02), self.N_gait, int(k == 0), max_iter, _p(xref), _p(fsteps),
                _p(self.std), _p(self.sti), _p(self.out), _p(self.info), _p(self.rho), _p(self.dbg), threads)
        return 0

    def gait(self):
        g = np.zeros((self.B, 4 * self.N_gait))
        for i in range(4):
            g[:, i * self.N_gait:(i + 1) * self.N_gait] = (self.sti[:, : self.N_gait] >> i) & 1
        return g

    def sgait(self):
        s = np.zeros((self.B, 192))
        for k in range(16):
            for a in range(12):
                s[:, 12 * k + a] = (self.sti[:, 32 + k] >> (a // 3)) & 1
        return s


def batched_mpc_class(emu):
    """A class with qrw_b200.BatchedMPC's constructor and reference-named methods, on the emulated kernel."""

    class EmuBatchedMPC(EmuMPC):
        def __init__(self, batch, dt=0.02, n_steps=16, T_gait=0.32, N_gait=20, device=0):
            assert dt == 0.02 and n_steps == 16
            super().__init__(emu, int(batch), int(N_gait))
            self.n = 16

        def get_latest_result(self, out=None):
            return self.out.copy()

        def get_gait(self):
            return self.gait()

        def get_Sgait(self):
            return self.sgait()

        def close(self):
            pass

    return EmuBatchedMPC
