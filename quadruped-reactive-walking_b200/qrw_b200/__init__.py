"""qrw_b200 -- host side of the B200-native batched controller core.

Thin Python over the C ABI of ``libqrw_b200.so`` (include/qrw_b200.h).  PyTorch / numpy appear only
at this boundary (device memory, streams); all arithmetic happens in the CUDA library.  There is
no CPU fallback: constructing any solver without the built library or without a CUDA device
raises ``QrwError``.
"""
from ._lib import QrwError, build_library, library_path, load_library  # noqa: F401
from .batched import BatchedInvKin, BatchedMPC, BatchedQPWBC, BatchedWbcTick, solo12_base_inertia, solo12_feet  # noqa: F401
