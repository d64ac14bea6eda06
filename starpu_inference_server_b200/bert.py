"""Host-side handle of one BERT weight replica (thin wrapper over the C-ABI).

Mirrors the role of one entry of ``InferenceParams::models.models_gpu``
(/root/reference/src/core/inference_params.hpp:46-51) created by
``clone_model_to_gpus`` (/root/reference/src/core/inference_runner.cpp:251-275): the
TorchScript state (or an HF-keyed state dict) is read ONCE at start-up, converted to
bf16 device buffers by ``spu_bert_replica_create``; afterwards no torch code runs on
the forward path.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import capi


@dataclass(frozen=True)
class BertConfig:
    layers: int = 12
    hidden: int = 768
    heads: int = 12
    ffn: int = 3072
    vocab: int = 30522
    max_pos: int = 512
    type_vocab: int = 2
    ln_eps: float = 1e-12


def _tensor_order(layers: int):
    names = ["embeddings.word_embeddings.weight", "embeddings.position_embeddings.weight",
             "embeddings.token_type_embeddings.weight", "embeddings.LayerNorm.weight",
             "embeddings.LayerNorm.bias"]
    for l in range(layers):
        p = f"encoder.layer.{l}."
        names += [p + s for s in (
            "attention.self.query.weight", "attention.self.query.bias",
            "attention.self.key.weight", "attention.self.key.bias",
            "attention.self.value.weight", "attention.self.value.bias",
            "attention.output.dense.weight", "attention.output.dense.bias",
            "attention.output.LayerNorm.weight", "attention.output.LayerNorm.bias",
            "intermediate.dense.weight", "intermediate.dense.bias",
            "output.dense.weight", "output.dense.bias",
            "output.LayerNorm.weight", "output.LayerNorm.bias")]
    return names


def config_from_state(state) -> BertConfig:
    """Infer the architecture from HF-keyed weights (heads = hidden/64)."""
    word = state["embeddings.word_embeddings.weight"]
    layers = 0
    while f"encoder.layer.{layers}.output.dense.weight" in state:
        layers += 1
    hidden = word.shape[1]
    return BertConfig(layers=layers, hidden=hidden, heads=hidden // 64,
                      ffn=state["encoder.layer.0.intermediate.dense.weight"].shape[0],
                      vocab=word.shape[0],
                      max_pos=state["embeddings.position_embeddings.weight"].shape[0],
                      type_vocab=state["embeddings.token_type_embeddings.weight"].shape[0])


class BertReplica:
    """One weight replica on one GPU + ``n_workspaces`` worker-private arenas."""

    def __init__(self, state, cfg: BertConfig | None = None, device: int = 0, max_batch: int = 32,
                 max_seq: int = 128, n_workspaces: int = 1, flags: int = capi.SPU_PREC_BF16):
        self.cfg = cfg or config_from_state(state)
        self.device, self.max_batch, self.max_seq = device, max_batch, max_seq
        names = _tensor_order(self.cfg.layers)
        keep = [np.ascontiguousarray(np.asarray(state[n], dtype=np.float32)) for n in names]
        arr = (C.c_void_p * len(keep))(*[a.ctypes.data for a in keep])
        desc = capi.BertDesc(self.cfg.layers, self.cfg.hidden, self.cfg.heads, self.cfg.ffn,
                             self.cfg.vocab, self.cfg.max_pos, self.cfg.type_vocab, self.cfg.ln_eps)
        h = C.c_void_p()
        capi.check(capi.lib().spu_bert_replica_create(C.byref(desc), arr, len(keep), device, max_batch,
                                                      max_seq, n_workspaces, flags, C.byref(h)))
        self._h = h

    @classmethod
    def from_handle(cls, handle, cfg: BertConfig, device: int, max_batch: int, max_seq: int):
        """Adopt a replica created elsewhere (the C++ TorchScript loader, loader.TorchScriptModule)."""
        self = cls.__new__(cls)
        self.cfg, self.device, self.max_batch, self.max_seq, self._h = cfg, device, max_batch, max_seq, handle
        return self

    @property
    def launches_per_forward(self) -> int:
        return capi.lib().spu_bert_forward_launch_count(self._h)

    def autotune(self, batch: int, seq: int, ws: int = 0, stream=None) -> int:
        """Warm-up autotune of the GEMM tiles for forwards of (batch, seq); start-up only.  Returns how many
        problems ended on another tile than the static cost model's."""
        changed = C.c_int32()
        capi.check(capi.lib().spu_bert_autotune(self._h, ws, batch, seq, capi.stream_ptr(stream), C.byref(changed)))
        return changed.value

    def tuned_configs(self):
        """[{N, K, epilogue, fp8, M, cta_group, block_n, us}] measured by autotune()."""
        n = capi.lib().spu_bert_tuned_configs(self._h, None, 0)
        buf = (C.c_int32 * (8 * max(n, 1)))()
        capi.lib().spu_bert_tuned_configs(self._h, buf, n)
        keys = ("N", "K", "epilogue", "fp8", "M", "cta_group", "block_n", "ns")
        rows = [dict(zip(keys, buf[8 * i:8 * i + 8])) for i in range(n)]
        for r in rows:
            r["us"] = r.pop("ns") / 1e3
        return rows

    def graph_count(self, ws: int = 0) -> int:
        return capi.lib().spu_bert_graph_count(self._h, ws)

    def forward_async(self, ids, mask, out, ws: int = 0, stream=None):
        """ids/mask: int64 CUDA tensors [B,S]; out: fp32 CUDA tensor [B,S,hidden]."""
        b, s = ids.shape
        capi.check(capi.lib().spu_bert_forward_async(self._h, ws, capi.ptr(ids), capi.ptr(mask), b, s,
                                                     capi.ptr(out), capi.stream_ptr(stream)))
        return out

    def forward_static_async(self, ids, mask, out, stream=None):
        """forward_async for a caller without a workspace: the activations go to the scratch array the library
        holds on every GPU (capi.SPU_STATIC_SCRATCH_BYTES; BERT-base: batch*seq <= 512), nothing is allocated.
        Calls on one GPU share that array: keep them ordered on the device."""
        b, s = ids.shape
        capi.check(capi.lib().spu_bert_forward_static_async(self._h, capi.ptr(ids), capi.ptr(mask), b, s,
                                                            capi.ptr(out), capi.stream_ptr(stream)))
        return out

    def forward_packed_async(self, ids, lens, out, ws: int = 0, stream=None, compact: bool = False):
        """Right-padded batch whose true lengths are known on the host: ids int64 CUDA [B,S], lens = B host ints.
        The encoder runs on sum(lens) tokens; out [B,S,hidden] gets zeros behind each length -- or, with
        `compact`, out is [sum(lens), hidden]: the packed rows only."""
        b, s = ids.shape
        lens = np.ascontiguousarray(np.asarray(lens, dtype=np.int32))
        assert lens.shape == (b,)
        fn = capi.lib().spu_bert_forward_packed_compact_async if compact else capi.lib().spu_bert_forward_packed_async
        capi.check(fn(self._h, ws, capi.ptr(ids), lens.ctypes.data, b, s, capi.ptr(out), capi.stream_ptr(stream)))
        return out

    def forward_requests_async(self, src_ids_table, src_mask_table, len_dev, n_req: int, seq: int, dst_table,
                               full_rows: bool = True, ws: int = 0, stream=None, status=None):
        """Stages 1-4 on per-request buffers (replica created with capi.SPU_REQUEST_IO).  The tables are int64 CUDA
        tensors of device-visible addresses (src_mask_table may be None), len_dev an int32 CUDA tensor.  `status`:
        optional int32 CUDA tensor of one element that receives 0, or 1 + a request row whose token ids leave the
        embedding table (spu_bert_forward_requests_checked_async)."""
        mask_tab = None if src_mask_table is None else capi.ptr(src_mask_table)
        if status is not None:
            capi.check(capi.lib().spu_bert_forward_requests_checked_async(
                self._h, ws, capi.ptr(src_ids_table), mask_tab, capi.ptr(len_dev), n_req, seq, capi.ptr(dst_table),
                1 if full_rows else 0, capi.ptr(status), capi.stream_ptr(stream)))
            return
        capi.check(capi.lib().spu_bert_forward_requests_async(
            self._h, ws, capi.ptr(src_ids_table), mask_tab,
            capi.ptr(len_dev), n_req, seq, capi.ptr(dst_table), 1 if full_rows else 0, capi.stream_ptr(stream)))

    def forward_profiled(self, ids, mask, out, ws: int = 0, stream=None):
        """Measurement hook: one forward with per-kernel CUDA events; returns
        {stage name: device ms summed over layers}.  Synchronises the stream."""
        b, s = ids.shape
        ms = (C.c_float * len(capi.STAGE_NAMES))()
        capi.check(capi.lib().spu_bert_forward_profiled(self._h, ws, capi.ptr(ids), capi.ptr(mask), b, s,
                                                        capi.ptr(out), capi.stream_ptr(stream), ms))
        return dict(zip(capi.STAGE_NAMES, list(ms)))

    def stage_chain_ms(self, ids, mask, out, stages=None, reps=5, ws: int = 0, stream=None) -> float:
        """Device ms per forward of only the named kernel classes (None = all), launched back to
        back in one CUDA graph and bracketed by CUDA events (measurement hook)."""
        b, s = ids.shape
        m = 0xFFFFFFFF if stages is None else sum(1 << capi.STAGE_NAMES.index(n) for n in stages)
        ms = C.c_float()
        capi.check(capi.lib().spu_bert_stage_chain_ms(self._h, ws, capi.ptr(ids), capi.ptr(mask), b, s, capi.ptr(out),
                                                      m, reps, capi.stream_ptr(stream), C.byref(ms)))
        return ms.value

    def calibrate_fp8(self, ids, mask, ws: int = 0, stream=None):
        """FP8 replicas: install per-tensor activation scales from this calibration batch."""
        b, s = ids.shape
        capi.check(capi.lib().spu_bert_calibrate_fp8(self._h, ws, capi.ptr(ids), capi.ptr(mask), b, s,
                                                     capi.stream_ptr(stream)))

    def set_fp8_mask(self, masks):
        """One byte per layer (OR of capi.FP8_GEMM_*): which GEMMs run on e4m3 operands.  Start-up only."""
        m = np.ascontiguousarray(np.asarray(masks, dtype=np.uint8))
        capi.check(capi.lib().spu_bert_set_fp8_mask(self._h, capi.ptr(m), len(m)))

    def fp8_mask(self):
        n = capi.lib().spu_bert_get_fp8_mask(self._h, None, 0)
        m = np.zeros(max(n, 1), dtype=np.uint8)
        capi.lib().spu_bert_get_fp8_mask(self._h, capi.ptr(m), n)
        return m[:n]

    def fp8_scales(self):
        n = capi.lib().spu_bert_fp8_scales(self._h, None, 0)
        buf = (C.c_float * max(n, 1))()
        capi.lib().spu_bert_fp8_scales(self._h, buf, n)
        return np.array(buf[:n], dtype=np.float32).reshape(-1, 8)

    def capture(self, ids, mask, out, ws: int = 0, stream=None):
        b, s = ids.shape
        capi.check(capi.lib().spu_bert_forward_capture(self._h, ws, capi.ptr(ids), capi.ptr(mask), b, s,
                                                       capi.ptr(out), capi.stream_ptr(stream)))

    def close(self):
        if getattr(self, "_h", None):
            capi.lib().spu_replica_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
