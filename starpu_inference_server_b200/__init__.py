"""B200-native replacement for the StarPU inference codelet hot path of
daxmawal/StarPU-Inference-Server.  The package holds only what the
path needs: ``csrc/`` (sm_100a kernels + the C-ABI of include/spu_b200.h), ``core/``
(the C++ host side mirroring the reference's codelet / slot-pool interface) and thin
ctypes plumbing for tests and the benchmark."""
from . import capi  # noqa: F401

__all__ = ["capi"]
