"""The drop-in boundary: libqrw_b200.so loads without a GPU, exports every symbol include/qrw_b200.h
declares, and fails loudly (error code + message, never a CPU fallback) when no CUDA device is usable."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, _have_gpu


def _declared():
    src = open(os.path.join(ROOT, "include", "qrw_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(qrw_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_are_exported_and_bound():
    from qrw_b200 import _lib

    _lib.build_library()
    L = C.CDLL(_lib.library_path())
    names = _declared()
    assert len(names) >= 25
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/qrw_b200.h but not exported"
    assert set(_lib.exported_symbols()) == set(names), "ctypes signatures out of sync with the header"


@pytest.mark.skipif(_have_gpu(), reason="checks the no-GPU behaviour")
def test_no_gpu_is_an_error_not_a_fallback():
    from qrw_b200 import BatchedMPC, BatchedQPWBC
    from qrw_b200._lib import QrwError

    with pytest.raises(QrwError, match="no CUDA device|no CPU fallback|CUDA"):
        BatchedMPC(4)
    with pytest.raises(QrwError):
        BatchedQPWBC(4)


def test_drop_in_module_surface():
    """Names and call signatures of python/gepadd.cpp:283-298 (the Boost.Python module being replaced)."""
    import inspect

    import libquadruped_reactive_walking as m

    assert m.add(4, 3) == 7 and m.sub(4, 3) == 1  # tests/python/test_add.py of the reference
    for cls, methods in {"MPC": ["run", "get_latest_result", "get_gait", "get_Sgait"],
                         "QPWBC": ["run", "compute", "get_f_res", "get_ddq_res", "get_H"],
                         "InvKin": ["refreshAndCompute", "get_q_step", "get_dq_cmd"], "Params": ["initialize"]}.items():
        for name in methods:
            assert callable(getattr(getattr(m, cls), name))
    assert list(inspect.signature(m.MPC.run).parameters)[1:] == ["num_iter", "xref_in", "fsteps_in"]
    p = m.Params()
    assert (p.dt_mpc, p.T_gait, p.N_gait, p.dt_wbc) == (0.02, 0.32, 20, 0.002)  # src/config_solo12.yaml
    assert np.isclose(p.T_mpc / p.dt_mpc, 16)
