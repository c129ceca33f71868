"""ctypes loader for libqrw_b200.so (built in-tree by __graft_entry__.build())."""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
_CSRC = os.path.join(_ROOT, "csrc")
_SO = os.path.join(_ROOT, "libqrw_b200.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-shared",
              "-Xcompiler", "-fPIC"]


class QrwError(RuntimeError):
    """An entry point of libqrw_b200.so returned non-zero (text from qrw_last_error())."""


def library_path():
    return _SO


def _stale():
    if not os.path.exists(_SO):
        return True
    t = os.path.getmtime(_SO)
    srcs = [os.path.join(_CSRC, f) for f in os.listdir(_CSRC)] + [os.path.join(os.path.dirname(_ROOT), "include", "qrw_b200.h")]
    return any(os.path.getmtime(s) > t for s in srcs if os.path.exists(s))


def build_library(force=False, verbose=False):
    """nvcc -gencode arch=compute_100a,code=sm_100a ... (cross-compiles without a GPU)."""
    if not force and not _stale():
        return _SO
    nvcc = os.environ.get("NVCC", "nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", _SO, os.path.join(_CSRC, "qrw_b200.cu")]
    subprocess.check_call(cmd)
    return _SO


_AB_SO = os.path.join(os.path.dirname(_ROOT), "tests", "ab", "libqrw_b200_ab.so")


def build_ab_library(force=False):
    """TEST INFRASTRUCTURE: the same source with -DQRW_AB_KERNELS -- the superseded MPC kernels (warp per robot, 24x24 block
    CTA) and the nested-dissection experiment next to the shipped one, selectable with qrw_mpc_set_kernel.  Only the A/B
    parity tests load it (tests/ab/); the product library contains one MPC kernel."""
    srcs = [os.path.join(_CSRC, f) for f in os.listdir(_CSRC)]
    if not force and os.path.exists(_AB_SO) and all(os.path.getmtime(x) <= os.path.getmtime(_AB_SO) for x in srcs):
        return _AB_SO
    os.makedirs(os.path.dirname(_AB_SO), exist_ok=True)
    nvcc = os.environ.get("NVCC", "nvcc")
    subprocess.check_call([nvcc] + NVCC_FLAGS + ["-DQRW_AB_KERNELS", "-o", _AB_SO, os.path.join(_CSRC, "qrw_b200.cu")])
    return _AB_SO


_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_vp = C.c_void_p
_lib = None

_SIGNATURES = {
    "qrw_last_error": (C.c_char_p, []),
    "qrw_launch_count": (C.c_longlong, []),
    "qrw_device_sm_count": (C.c_int, [C.c_int, _ip]),
    "qrw_fp64_peak": (C.c_int, [C.c_int, C.POINTER(C.c_double)]),
    "qrw_mpc_create": (C.c_int, [C.POINTER(_vp), C.c_int, C.c_double, C.c_int, C.c_double, C.c_int, C.c_int]),
    "qrw_mpc_destroy": (C.c_int, [_vp]),
    "qrw_mpc_run_device": (C.c_int, [_vp, C.c_int, _vp, _vp, _vp]),
    "qrw_mpc_run_host": (C.c_int, [_vp, C.c_int, _vp, _vp]),
    "qrw_mpc_run_indexed_device": (C.c_int, [_vp, _vp, _vp, _vp, _vp]),
    "qrw_mpc_server_create": (C.c_int, [C.POINTER(_vp), _vp]),
    "qrw_mpc_server_destroy": (C.c_int, [_vp]),
    "qrw_mpc_server_slot_doubles": (C.c_int, [_vp, _ip, _ip]),
    "qrw_mpc_server_poll_host": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _ip]),
    "qrw_mpc_server_loop_host": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, C.c_int, C.c_longlong, C.POINTER(C.c_longlong)]),
    "qrw_mpc_result_host": (C.c_int, [_vp, _vp]),
    "qrw_mpc_result_device": (C.c_int, [_vp, C.POINTER(_vp)]),
    "qrw_mpc_gait_host": (C.c_int, [_vp, _vp]),
    "qrw_mpc_sgait_host": (C.c_int, [_vp, _vp]),
    "qrw_mpc_info_host": (C.c_int, [_vp, _vp, _vp, _vp, _vp]),
    "qrw_mpc_set_max_iter": (C.c_int, [_vp, C.c_int]),
    "qrw_mpc_debug_host": (C.c_int, [_vp, _vp]),
    "qrw_mpc_set_kernel": (C.c_int, [_vp, C.c_int]),
    "qrw_mpc_set_work_order": (C.c_int, [_vp, C.c_int]),
    "qrw_mpc_state_bytes": (C.c_int, [_vp, C.POINTER(C.c_size_t)]),
    "qrw_mpc_state_save": (C.c_int, [_vp, _vp, _vp]),
    "qrw_mpc_state_load": (C.c_int, [_vp, _vp, _vp]),
    "qrw_mpc_pinned": (C.c_int, [_vp, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp)]),
    "qrw_wbc_create": (C.c_int, [C.POINTER(_vp), C.c_int, C.c_int]),
    "qrw_wbc_destroy": (C.c_int, [_vp]),
    "qrw_wbc_run_device": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "qrw_wbc_run_host": (C.c_int, [_vp, _vp, _vp, _vp, _vp]),
    "qrw_wbc_result_host": (C.c_int, [_vp, _vp, _vp, _vp]),
    "qrw_wbc_result_device": (C.c_int, [_vp, C.POINTER(_vp), C.POINTER(_vp)]),
    "qrw_wbc_info_host": (C.c_int, [_vp, _vp, _vp, _vp]),
    "qrw_invkin_run_device": (C.c_int, [C.c_int] + [_vp] * 13),
    "qrw_invkin_run_host": (C.c_int, [C.c_int] + [_vp] * 12),
    "qrw_wbctick_create": (C.c_int, [C.POINTER(_vp), C.c_int, C.c_int]),
    "qrw_wbctick_destroy": (C.c_int, [_vp]),
    "qrw_wbctick_run_device": (C.c_int, [_vp] * 9),
    "qrw_wbctick_run_fstride_device": (C.c_int, [_vp, _vp, _vp, _vp, C.c_int, _vp, _vp, _vp, _vp, _vp]),
    "qrw_wbctick_run_host": (C.c_int, [_vp] * 8),
    "qrw_wbctick_result_host": (C.c_int, [_vp] * 6),
    "qrw_wbctick_result_device": (C.c_int, [_vp] + [C.POINTER(_vp)] * 4),
    "qrw_wbctick_debug_host": (C.c_int, [_vp] * 4),
    "qrw_wbctick_set_debug": (C.c_int, [_vp, C.c_int]),
    "qrw_wbctick_set_group": (C.c_int, [_vp, C.c_int]),
    "qrw_wbctick_noalloc_device": (C.c_int, [C.c_int, C.c_int] + [_vp] * 11 + [_vp]),
    "qrw_wbctick_set_desired_host": (C.c_int, [_vp, _vp, _vp]),
    "qrw_wbctick_run_ctrl_device": (C.c_int, [_vp, C.c_double, _vp, _vp, C.c_int, _vp, _vp, _vp, _vp, _vp]),
    "qrw_solo12_base_inertia": (C.c_int, [_vp]),
    "qrw_solo12_feet_host": (C.c_int, [C.c_int] + [_vp] * 7),
    "qrw_planner_create": (C.c_int, [C.POINTER(_vp), C.c_int, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int, C.c_int,
                                     C.c_double, C.c_int]),
    "qrw_planner_destroy": (C.c_int, [_vp]),
    "qrw_planner_configure": (C.c_int, [_vp, _vp, _vp, _vp, C.c_double, C.c_double]),
    "qrw_planner_run_device": (C.c_int, [_vp, C.c_int, C.c_int, _vp, _vp, _vp, _vp, _vp]),
    "qrw_planner_run_host": (C.c_int, [_vp, C.c_int, C.c_int, _vp, _vp, _vp, _vp]),
    "qrw_planner_outputs_device": (C.c_int, [_vp] + [C.POINTER(_vp)] * 10),
    "qrw_planner_outputs_host": (C.c_int, [_vp] * 11),
    "qrw_planner_state_host": (C.c_int, [_vp] * 7),
    "qrw_planner_set_target_host": (C.c_int, [_vp, _vp]),
    "qrw_tick_create": (C.c_int, [C.POINTER(_vp), C.c_int, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int, C.c_double, _vp,
                                  C.c_int]),
    "qrw_tick_destroy": (C.c_int, [_vp]),
    "qrw_tick_run_device": (C.c_int, [_vp, C.c_int, _vp, _vp, _vp, _vp, _vp]),
    "qrw_tick_run_host": (C.c_int, [_vp, C.c_int, _vp, _vp, _vp, _vp]),
    "qrw_tick_result_host": (C.c_int, [_vp] * 5),
    "qrw_tick_result_device": (C.c_int, [_vp] + [C.POINTER(_vp)] * 4),
    "qrw_tick_parts": (C.c_int, [_vp] + [C.POINTER(_vp)] * 3),
    "qrw_tick_capture_period": (C.c_int, [_vp]),
    "qrw_tick_period_buffers": (C.c_int, [_vp] + [C.POINTER(_vp)] * 7),
    "qrw_tick_replay_period": (C.c_int, [_vp, _vp]),
    "qrw_tick_period_host": (C.c_int, [_vp] * 8),
}


def exported_symbols():
    """Names the ctypes layer binds -- tests compare this list with include/qrw_b200.h."""
    return sorted(_SIGNATURES)


def _bind(path):
    L = C.CDLL(path)
    for name, (res, args) in _SIGNATURES.items():
        f = getattr(L, name)
        f.restype = res
        f.argtypes = args
    return L


_others = {}


def load_library(path=None):
    """Load libqrw_b200.so; raises QrwError (never falls back) if it has not been built.  `path` loads another build of the
    same ABI next to it (the A/B test library) without replacing the default."""
    global _lib
    if path is not None:
        if path not in _others:
            if not os.path.exists(path):
                raise QrwError(f"{path} not found")
            _others[path] = _bind(path)
        return _others[path]
    if _lib is None:
        if not os.path.exists(_SO):
            raise QrwError(
                f"{_SO} not found: build the CUDA extension first (python -c 'import __graft_entry__ as g; g.build()'). "
                "qrw_b200 has no CPU fallback.")
        _lib = _bind(_SO)
    return _lib


def check(rc):
    if rc != 0:
        raise QrwError(load_library().qrw_last_error().decode() or f"error code {rc}")
