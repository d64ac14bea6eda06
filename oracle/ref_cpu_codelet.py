"""ctypes driver of the reference's CPU codelet body (TEST INFRASTRUCTURE ONLY, see oracle/__init__.py).

``oracle/_ref/libref_cpu_codelet.so`` = the reference's own ``src/core/tensor_builder.cpp`` (compiled from where it
lies under /root/reference by ``make -C oracle _ref``) + the glue of ``oracle/ref_cpu_codelet.cpp`` restating
``run_inference`` / ``cpu_inference_func`` (/root/reference/src/core/starpu_setup.cpp:594-624, 784-801), linked against
the torch wheel's LibTorch.  The TorchScript artifact it loads stands in for ``models/bert_libtorch.pt`` /
``models/vit_l_16.pt``: the traced oracle module (HF BertModel does not trace under transformers 5.5) and
``torch.jit.script(torchvision vit)`` as /root/reference/models/import_vit.py:51-55 exports it.

Used by bench.py (``cpu_baseline`` leg and ``--impl reference``: kind "reference") and by the tests that pin the
Python oracle against what LibTorch computes through the reference's TensorBuilder.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_ref", "libref_cpu_codelet.so")
_DT = {np.dtype(np.float32): 0, np.dtype(np.int64): 3}  # include/spu_host.h SPU_DT_*


def available() -> bool:
    return os.path.exists(LIB_PATH)


_lib = None


def lib():
    global _lib
    if _lib is None:
        import torch  # noqa: F401  (loads libtorch / libc10 the library depends on)

        L = C.CDLL(LIB_PATH)
        L.ref_cpu_codelet_open.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_void_p)]
        L.ref_cpu_codelet_open.restype = C.c_int
        L.ref_cpu_codelet_close.argtypes = [C.c_void_p]
        L.ref_cpu_codelet_close.restype = None
        L.ref_cpu_codelet_run.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_int),
                                          C.POINTER(C.c_longlong), C.POINTER(C.c_int), C.c_void_p, C.c_ulonglong,
                                          C.c_int, C.c_int, C.POINTER(C.c_double)]
        L.ref_cpu_codelet_run.restype = C.c_int
        L.ref_cpu_codelet_last_error.restype = C.c_char_p
        _lib = L
    return _lib


def export_bert_torchscript(w, spec, path, batch=1, seq=128):
    """Trace the oracle module the way models/import_bert-base-uncased.py:8-25 traces its wrapper."""
    import torch

    from . import bert_oracle as bo

    ids, mask = bo.make_inputs(spec, batch, seq, seed=1)
    m = bo.BertOracleModule(w, spec).eval()
    with torch.no_grad():
        ts = torch.jit.trace(m, (torch.from_numpy(ids), torch.from_numpy(mask)))
    ts.save(path)
    return path


def export_vit_torchscript(w, spec, path):
    """torch.jit.script of torchvision's VisionTransformer (models/import_vit.py:51-55)."""
    import torch

    from . import vit_oracle as vo

    torch.jit.script(vo.torchvision_model(w, spec).eval()).save(path)
    return path


class RefCpuCodelet:
    """One loaded TorchScript module behind the reference's TensorBuilder -> forward -> copy_output_to_buffer."""

    def __init__(self, path: str, threads: int | None = None):
        h = C.c_void_p()
        if lib().ref_cpu_codelet_open(path.encode(), int(threads or 0), C.byref(h)) != 0:
            raise RuntimeError(lib().ref_cpu_codelet_last_error().decode(errors="replace"))
        self._h = h

    def run(self, inputs, out: np.ndarray, warmup: int = 0, iters: int = 1):
        """inputs: C-contiguous numpy arrays (int64 / fp32), out: preallocated output array the codelet copies into.
        Returns the per-task wall times in ms (length `iters`)."""
        inputs = [np.ascontiguousarray(a) for a in inputs]
        ptrs = (C.c_void_p * len(inputs))(*[a.ctypes.data for a in inputs])
        dts = (C.c_int * len(inputs))(*[_DT[a.dtype] for a in inputs])
        flat = [d for a in inputs for d in a.shape]
        dims = (C.c_longlong * len(flat))(*flat)
        nd = (C.c_int * len(inputs))(*[a.ndim for a in inputs])
        ms = (C.c_double * max(iters, 1))()
        assert out.flags["C_CONTIGUOUS"]
        if lib().ref_cpu_codelet_run(self._h, len(inputs), ptrs, dts, dims, nd, out.ctypes.data, out.nbytes, warmup,
                                     iters, ms) != 0:
            raise RuntimeError(lib().ref_cpu_codelet_last_error().decode(errors="replace"))
        return [ms[i] for i in range(iters)]

    def close(self):
        if getattr(self, "_h", None):
            lib().ref_cpu_codelet_close(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
