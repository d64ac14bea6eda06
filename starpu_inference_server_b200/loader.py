"""ctypes binding of include/spu_loader.h (libspu_loader.so): the C++ TorchScript weight loader.

Start-up only.  The library links the wheel's LibTorch; nothing on the forward path does."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import capi

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libspu_loader.so")
KIND_BERT, KIND_VIT = 0, 1
SYMBOLS = ["spu_loader_open", "spu_loader_open_ex", "spu_loader_close", "spu_loader_kind", "spu_loader_num_tensors", "spu_loader_tensor",
           "spu_loader_bert_desc", "spu_loader_vit_desc", "spu_loader_create_replica", "spu_loader_last_error"]

_lib = None


class LoaderError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run `python -m starpu_inference_server_b200.build`")
    capi.lib()  # libspu_b200.so first (the loader's DT_NEEDED resolves through $ORIGIN as well)
    L = C.CDLL(LIB_PATH)
    vp, i32 = C.c_void_p, C.c_int32
    L.spu_loader_open.argtypes = [C.c_char_p, i32, C.POINTER(vp)]
    L.spu_loader_open_ex.argtypes = [C.c_char_p, i32, i32, C.c_float, C.POINTER(vp)]
    L.spu_loader_close.argtypes = [vp]
    L.spu_loader_close.restype = None
    L.spu_loader_kind.argtypes = [vp]
    L.spu_loader_num_tensors.argtypes = [vp]
    L.spu_loader_tensor.argtypes = [vp, i32, C.POINTER(vp), C.POINTER(C.c_uint64), C.c_char_p, C.c_size_t]
    L.spu_loader_bert_desc.argtypes = [vp, C.POINTER(capi.BertDesc)]
    L.spu_loader_vit_desc.argtypes = [vp, C.POINTER(capi.VitDesc)]
    L.spu_loader_create_replica.argtypes = [vp, i32, i32, i32, i32, C.c_uint32, C.POINTER(vp)]
    L.spu_loader_last_error.restype = C.c_char_p
    _lib = L
    return L


class TorchScriptModule:
    """`torch::jit::load` + `eval()` of a served artifact, done in C++ (load_model, inference_runner.cpp:243-249)."""

    def __init__(self, path: str, kind: int, heads: int = 0, ln_eps: float = 0.0):
        """heads / ln_eps: 0 = read them from the artifact (the loader refuses rather than guess); a positive value,
        e.g. from models/*.yml, overrides."""
        h = C.c_void_p()
        if lib().spu_loader_open_ex(os.fsencode(path), kind, heads, ln_eps, C.byref(h)) != 0:
            raise LoaderError(lib().spu_loader_last_error().decode(errors="replace"))
        self._h, self.kind = h, kind

    def tensors(self):
        """[(name, fp32 numpy copy)] in replica-create order."""
        out = []
        for i in range(lib().spu_loader_num_tensors(self._h)):
            p, n, name = C.c_void_p(), C.c_uint64(), C.create_string_buffer(256)
            assert lib().spu_loader_tensor(self._h, i, C.byref(p), C.byref(n), name, 256) == 0
            a = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=(n.value,)).copy()
            out.append((name.value.decode(), a))
        return out

    def desc(self):
        d = capi.BertDesc() if self.kind == KIND_BERT else capi.VitDesc()
        fn = lib().spu_loader_bert_desc if self.kind == KIND_BERT else lib().spu_loader_vit_desc
        if fn(self._h, C.byref(d)) != 0:
            raise LoaderError(lib().spu_loader_last_error().decode(errors="replace"))
        return d

    def create_replica(self, device=0, max_batch=32, max_seq=128, n_workspaces=1, flags=capi.SPU_PREC_BF16):
        """clone_model_to_gpus for one device: returns a BertReplica / VitReplica owning the new handle."""
        from .bert import BertConfig, BertReplica
        from .vit import VitConfig, VitReplica

        h = C.c_void_p()
        st = lib().spu_loader_create_replica(self._h, device, max_batch, max_seq, n_workspaces, flags, C.byref(h))
        if st != 0:
            raise capi.SpuError(st, lib().spu_loader_last_error().decode(errors="replace"))
        d = self.desc()
        if self.kind == KIND_BERT:
            cfg = BertConfig(d.layers, d.hidden, d.heads, d.ffn, d.vocab, d.max_pos, d.type_vocab, d.ln_eps)
            return BertReplica.from_handle(h, cfg, device, max_batch, max_seq)
        cfg = VitConfig(d.layers, d.hidden, d.heads, d.mlp, d.patch, d.image, d.channels, d.classes, d.ln_eps)
        return VitReplica.from_handle(h, cfg, device, max_batch)

    def close(self):
        if getattr(self, "_h", None):
            lib().spu_loader_close(self._h)
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
