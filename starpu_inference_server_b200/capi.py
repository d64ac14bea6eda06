"""ctypes binding of include/spu_b200.h (libspu_b200.so).

This is plumbing for tests/bench: PyTorch supplies device memory and streams, every
computation goes through the C-ABI.  There is no fallback: a missing library or a
non-sm_100 device raises.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libspu_b200.so")

SPU_OK = 0
SPU_ERR_INVALID, SPU_ERR_CUDA, SPU_ERR_CAPACITY, SPU_ERR_UNSUPPORTED, SPU_ERR_NOMEM = 1, 2, 3, 4, 5
SPU_EPI_BIAS, SPU_EPI_BIAS_GELU, SPU_EPI_BIAS_RESID = 0, 1, 2
SPU_PREC_BF16, SPU_PREC_FP8_E4M3 = 0, 1
SPU_FORWARD_AUTO_GRAPH = 1 << 16  # capture a CUDA graph on the second sighting of a pointer tuple
SPU_OUTPUT_BF16 = 1 << 18  # output0 is bf16 [B,S,H] (response path)
SPU_LN_FUSED = 1 << 19  # residual + LayerNorm in the N = hidden GEMM epilogues
SPU_REQUEST_IO = 1 << 17  # allocate the staging spu_bert_forward_requests_async needs
SPU_STATIC_SCRATCH_BYTES = 9 << 20  # per-GPU scratch array of spu_bert_forward_static_async
FP8_GEMM_QKV, FP8_GEMM_OUT, FP8_GEMM_FFN1, FP8_GEMM_FFN2, FP8_GEMM_ALL = 1, 2, 4, 8, 15
STAGE_NAMES = ["embed_ln", "gemm_qkv", "attention", "gemm_attn_out", "ln1", "gemm_ffn1", "gemm_ffn2", "ln2"]

# every symbol include/spu_b200.h declares (checked by tests/test_capi_symbols.py)
SYMBOLS = [
    "spu_bert_replica_create", "spu_replica_destroy", "spu_bert_forward_async", "spu_bert_forward_static_async",
    "spu_bert_forward_launch_count", "spu_bert_forward_requests_async", "spu_bert_forward_requests_checked_async", "spu_bert_forward_capture", "spu_bert_forward_profiled", "spu_gather_pad_mask_async",
    "spu_gather_segments_async", "spu_scatter_segments_async", "spu_gemm_bf16_async", "spu_gemm_bf16_cfg_async",
    "spu_attention_async", "spu_attention_cfg_async", "spu_layernorm_async", "spu_bert_embed_async", "spu_affine_f32_async",
    "spu_last_error", "spu_abi_version", "spu_device_count", "spu_device_is_sm100",
    "spu_bert_output_bytes", "spu_vit_replica_create", "spu_vit_forward_async", "spu_vit_input_bytes",
    "spu_vit_output_bytes", "spu_bert_stage_chain_ms", "spu_bert_calibrate_fp8", "spu_bert_fp8_scales",
    "spu_gemm_fp8_async", "spu_quantize_e4m3_async", "spu_bert_set_fp8_mask", "spu_bert_get_fp8_mask",
    "spu_gemm_ln_bf16_async", "spu_bert_graph_count", "spu_bert_autotune", "spu_bert_tuned_configs",
    "spu_bert_forward_packed_async", "spu_bert_packed_supported", "spu_bert_vocab_size",
    "spu_bert_forward_packed_compact_async", "spu_vit_gemm_chain_ms",
]


class SpuError(RuntimeError):
    def __init__(self, status, msg):
        super().__init__(f"spu status {status}: {msg}")
        self.status = status


class BertDesc(C.Structure):
    _fields_ = [("layers", C.c_int32), ("hidden", C.c_int32), ("heads", C.c_int32),
                ("ffn", C.c_int32), ("vocab", C.c_int32), ("max_pos", C.c_int32),
                ("type_vocab", C.c_int32), ("ln_eps", C.c_float)]


class VitDesc(C.Structure):
    _fields_ = [("layers", C.c_int32), ("hidden", C.c_int32), ("heads", C.c_int32), ("mlp", C.c_int32),
                ("patch", C.c_int32), ("image", C.c_int32), ("channels", C.c_int32), ("classes", C.c_int32),
                ("ln_eps", C.c_float)]


_lib = None


def lib():
    """Load libspu_b200.so (built in-tree by starpu_inference_server_b200.build)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: run `python -m starpu_inference_server_b200.build` "
            "(there is no CPU or PyTorch fallback for this path)")
    L = C.CDLL(LIB_PATH)
    vp, i32, i64, u64, f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_float
    L.spu_bert_replica_create.argtypes = [C.POINTER(BertDesc), C.POINTER(vp), i32, i32, i32, i32, i32,
                                          C.c_uint32, C.POINTER(vp)]
    L.spu_replica_destroy.argtypes = [vp]
    L.spu_bert_forward_async.argtypes = [vp, i32, vp, vp, i32, i32, vp, vp]
    L.spu_bert_forward_capture.argtypes = [vp, i32, vp, vp, i32, i32, vp, vp]
    L.spu_bert_forward_static_async.argtypes = [vp, vp, vp, i32, i32, vp, vp]
    L.spu_bert_forward_static_async.restype = i32
    L.spu_bert_forward_profiled.argtypes = [vp, i32, vp, vp, i32, i32, vp, vp, C.POINTER(f32)]
    L.spu_bert_forward_packed_async.argtypes = [vp, i32, vp, vp, i32, i32, vp, vp]
    L.spu_bert_forward_packed_async.restype = i32
    L.spu_bert_forward_packed_compact_async.argtypes = [vp, i32, vp, vp, i32, i32, vp, vp]
    L.spu_bert_forward_packed_compact_async.restype = i32
    L.spu_vit_gemm_chain_ms.argtypes = [vp, i32, vp, i32, vp, i32, vp, C.POINTER(f32)]
    L.spu_vit_gemm_chain_ms.restype = i32
    L.spu_bert_vocab_size.argtypes = [vp]
    L.spu_bert_vocab_size.restype = i32
    L.spu_bert_packed_supported.argtypes = [vp, i32]
    L.spu_bert_packed_supported.restype = i32
    L.spu_bert_autotune.argtypes = [vp, i32, i32, i32, vp, C.POINTER(i32)]
    L.spu_bert_autotune.restype = i32
    L.spu_bert_tuned_configs.argtypes = [vp, C.POINTER(i32), i32]
    L.spu_bert_tuned_configs.restype = i32
    L.spu_bert_graph_count.argtypes = [vp, i32]
    L.spu_bert_graph_count.restype = i32
    L.spu_bert_forward_launch_count.argtypes = [vp]
    L.spu_gather_pad_mask_async.argtypes = [vp, vp, vp, i32, i32, i64, vp, vp, vp]
    L.spu_gather_segments_async.argtypes = [vp, vp, vp, i32, vp, vp]
    L.spu_scatter_segments_async.argtypes = [vp, vp, vp, vp, i32, vp]
    L.spu_gemm_bf16_async.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, vp]
    L.spu_gemm_ln_bf16_async.argtypes = [vp, vp, vp, vp, vp, vp, f32, vp, vp, i32, i32, i32, i32, i32, vp]
    L.spu_gemm_bf16_cfg_async.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp]
    L.spu_attention_async.argtypes = [vp, vp, vp, i32, i32, i32, vp]
    L.spu_bert_forward_requests_async.argtypes = [vp, i32, vp, vp, vp, i32, i32, vp, i32, vp]
    L.spu_bert_forward_requests_checked_async.argtypes = [vp, i32, vp, vp, vp, i32, i32, vp, i32, vp, vp]
    L.spu_bert_forward_requests_checked_async.restype = i32
    L.spu_attention_cfg_async.argtypes = [vp, vp, vp, i32, i32, i32, i32, i32, f32, vp]
    L.spu_layernorm_async.argtypes = [vp, vp, vp, f32, vp, vp, i32, i32, vp]
    L.spu_bert_embed_async.argtypes = [vp, vp, vp, vp, vp, vp, vp, f32, i32, vp, vp, vp, i32, i32, i32, vp]
    L.spu_affine_f32_async.argtypes = [vp, vp, u64, f32, f32, vp]
    L.spu_vit_replica_create.argtypes = [C.POINTER(VitDesc), C.POINTER(vp), i32, i32, i32, i32, C.c_uint32,
                                         C.POINTER(vp)]
    L.spu_vit_forward_async.argtypes = [vp, i32, vp, i32, vp, vp]
    L.spu_bert_stage_chain_ms.argtypes = [vp, i32, vp, vp, i32, i32, vp, C.c_uint32, i32, vp, C.POINTER(f32)]
    L.spu_bert_calibrate_fp8.argtypes = [vp, i32, vp, vp, i32, i32, vp]
    L.spu_bert_fp8_scales.argtypes = [vp, C.POINTER(f32), i32]
    L.spu_bert_set_fp8_mask.argtypes = [vp, vp, i32]
    L.spu_bert_get_fp8_mask.argtypes = [vp, vp, i32]
    L.spu_gemm_fp8_async.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, f32, f32, vp]
    L.spu_quantize_e4m3_async.argtypes = [vp, vp, u64, f32, vp]
    if hasattr(L, "spu_debug_gemm_trace"):  # only in SPU_DEBUG_HOOKS builds (include/spu_b200_debug.h)
        L.spu_debug_gemm_trace.argtypes = [vp, i32]
        L.spu_debug_gemm_trace.restype = None
        L.spu_debug_gemm_mode.argtypes = [i32]
        L.spu_debug_gemm_mode.restype = None
    L.spu_bert_output_bytes.argtypes = [vp, i32, i32]
    L.spu_vit_input_bytes.argtypes = [vp, i32]
    L.spu_vit_output_bytes.argtypes = [vp, i32]
    L.spu_last_error.restype = C.c_char_p
    for name in SYMBOLS:
        fn = getattr(L, name)
        if name.endswith("_bytes"):
            fn.restype = u64
        elif name not in ("spu_last_error",):
            fn.restype = i32
    _lib = L
    return L


def debug_hooks():
    """The measurement hooks of include/spu_b200_debug.h; raises unless the library was built with them."""
    L = lib()
    if not hasattr(L, "spu_debug_gemm_trace"):
        raise RuntimeError("libspu_b200.so was built without -DSPU_DEBUG_HOOKS: rebuild with "
                           "`SPU_DEBUG_HOOKS=1 python -m starpu_inference_server_b200.build --force`")
    return L


def check(status):
    if status != SPU_OK:
        raise SpuError(status, lib().spu_last_error().decode(errors="replace"))


def ptr(t):
    """Device/host address of a torch tensor or numpy array (None -> NULL)."""
    if t is None:
        return None
    if hasattr(t, "data_ptr"):
        return C.c_void_p(t.data_ptr())
    return C.c_void_p(t.ctypes.data)


def stream_ptr(stream=None):
    import torch

    s = stream if stream is not None else torch.cuda.current_stream()
    return C.c_void_p(s.cuda_stream)
