"""Host-side handle of one ViT weight replica (thin wrapper over the C-ABI).

Same role as ``bert.BertReplica``: the scripted torchvision module
(/root/reference/models/import_vit.py:51-55) is read ONCE at start-up, its fp32 state is
handed to ``spu_vit_replica_create``; afterwards no torch code runs on the forward path.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import capi


@dataclass(frozen=True)
class VitConfig:
    layers: int = 24
    hidden: int = 1024
    heads: int = 16
    mlp: int = 4096
    patch: int = 16
    image: int = 224
    channels: int = 3
    classes: int = 1000
    ln_eps: float = 1e-6

    @property
    def seq(self) -> int:
        return (self.image // self.patch) ** 2 + 1


def _tensor_order(layers: int):
    names = ["conv_proj.weight", "conv_proj.bias", "class_token", "encoder.pos_embedding"]
    for l in range(layers):
        p = f"encoder.layers.encoder_layer_{l}."
        names += [p + s for s in (
            "ln_1.weight", "ln_1.bias", "self_attention.in_proj_weight", "self_attention.in_proj_bias",
            "self_attention.out_proj.weight", "self_attention.out_proj.bias", "ln_2.weight", "ln_2.bias",
            "mlp.0.weight", "mlp.0.bias", "mlp.3.weight", "mlp.3.bias")]
    return names + ["encoder.ln.weight", "encoder.ln.bias", "heads.head.weight", "heads.head.bias"]


def config_from_state(state, image: int = 224) -> VitConfig:
    conv = state["conv_proj.weight"]
    layers = 0
    while f"encoder.layers.encoder_layer_{layers}.ln_1.weight" in state:
        layers += 1
    hidden, channels, patch = conv.shape[0], conv.shape[1], conv.shape[2]
    n_tok = state["encoder.pos_embedding"].shape[1]
    side = int(round((n_tok - 1) ** 0.5))
    return VitConfig(layers=layers, hidden=hidden, heads=hidden // 64,
                     mlp=state["encoder.layers.encoder_layer_0.mlp.0.weight"].shape[0], patch=patch,
                     image=side * patch, channels=channels, classes=state["heads.head.weight"].shape[0])


class VitReplica:
    """One ViT weight replica on one GPU + ``n_workspaces`` worker-private arenas."""

    def __init__(self, state, cfg: VitConfig | None = None, device: int = 0, max_batch: int = 64,
                 n_workspaces: int = 1, flags: int = capi.SPU_PREC_BF16):
        self.cfg = cfg or config_from_state(state)
        self.device, self.max_batch = device, max_batch
        keep = [np.ascontiguousarray(np.asarray(state[n], dtype=np.float32)) for n in _tensor_order(self.cfg.layers)]
        arr = (C.c_void_p * len(keep))(*[a.ctypes.data for a in keep])
        c = self.cfg
        desc = capi.VitDesc(c.layers, c.hidden, c.heads, c.mlp, c.patch, c.image, c.channels, c.classes, c.ln_eps)
        h = C.c_void_p()
        capi.check(capi.lib().spu_vit_replica_create(C.byref(desc), arr, len(keep), device, max_batch, n_workspaces,
                                                     flags, C.byref(h)))
        self._h = h

    @classmethod
    def from_handle(cls, handle, cfg: VitConfig, device: int, max_batch: int):
        """Adopt a replica created elsewhere (the C++ TorchScript loader, loader.TorchScriptModule)."""
        self = cls.__new__(cls)
        self.cfg, self.device, self.max_batch, self._h = cfg, device, max_batch, handle
        return self

    @property
    def launches_per_forward(self) -> int:
        return 3 + 7 * self.cfg.layers - 1 + 3

    def forward_async(self, pixels, logits, ws: int = 0, stream=None):
        """pixels: fp32 CUDA tensor [B,C,H,W]; logits: fp32 CUDA tensor [B,classes]."""
        capi.check(capi.lib().spu_vit_forward_async(self._h, ws, capi.ptr(pixels), pixels.shape[0], capi.ptr(logits),
                                                    capi.stream_ptr(stream)))
        return logits

    def gemm_chain_ms(self, pixels, logits, reps: int = 3, ws: int = 0, stream=None) -> float:
        """Device ms per forward of the dense GEMMs alone (measurement hook; needs a non-default stream)."""
        ms = C.c_float()
        capi.check(capi.lib().spu_vit_gemm_chain_ms(self._h, ws, capi.ptr(pixels), pixels.shape[0], capi.ptr(logits), reps,
                                                    capi.stream_ptr(stream), C.byref(ms)))
        return ms.value

    def close(self):
        if getattr(self, "_h", None):
            capi.lib().spu_replica_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
