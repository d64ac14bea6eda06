"""ctypes binding of include/spu_host.h (libspu_host.so): the C++ host side -- StarPU-shaped
codelet, slot pools, batch assembly, result slicing -- driven from Python for tests and
bench.py's end-to-end leg.  Plumbing only; there is no fallback if the library is missing.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import capi

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libspu_host.so")
MAX_IO, MAX_DIMS, MAX_DEV = 16, 8, 32
DT = {"f32": 0, "f64": 1, "i32": 2, "i64": 3, "i8": 4, "u8": 5, "f16": 6, "bf16": 7, "bool": 8}
NP_DT = {0: np.float32, 1: np.float64, 2: np.int32, 3: np.int64, 4: np.int8, 5: np.uint8, 6: np.float16}
MODEL_BERT, MODEL_VIT, MODEL_AFFINE = 0, 1, 2

SYMBOLS = [
    "spu_server_create", "spu_server_destroy", "spu_server_submit_batch", "spu_server_enqueue", "spu_server_wait",
    "spu_server_wait_all", "spu_server_stats", "spu_host_last_error", "spu_host_codelet_fields",
    "spu_host_buffer_byte_size", "spu_host_select_replica", "spu_host_gather_to_slot", "spu_host_slice_rows",
    "spu_host_can_merge", "spu_host_call_cuda_func_affine", "spu_host_can_merge_varlen", "spu_host_gather_padded",
    "spu_host_replica_assignments", "spu_host_nvtx_capture", "spu_server_offered_load",
    "spu_host_tile_aware_wants_more", "spu_host_request_table", "spu_host_masks_all_ones",
]


class TensorCfg(C.Structure):
    _fields_ = [("name", C.c_char * 64), ("dtype", C.c_int32), ("ndims", C.c_int32), ("dims", C.c_int64 * MAX_DIMS)]


class ServerCfg(C.Structure):
    _fields_ = [("n_devices", C.c_int32), ("device_ids", C.c_int32 * MAX_DEV), ("nworker_per_cuda", C.c_int32),
                ("cuda_pipeline", C.c_int32), ("pool_size", C.c_int32), ("max_batch_size", C.c_int32),
                ("batch_coalesce_timeout_ms", C.c_int32), ("n_inputs", C.c_int32), ("n_outputs", C.c_int32),
                ("inputs", TensorCfg * MAX_IO), ("outputs", TensorCfg * MAX_IO), ("model_kind", C.c_int32),
                ("replicas", C.c_void_p * MAX_DEV), ("n_workspaces", C.c_int32), ("affine_scale", C.c_float),
                ("affine_shift", C.c_float), ("variable_length", C.c_int32),
                ("length_bucket", C.c_int32), ("gpu_model_replication", C.c_int32), ("tile_rows", C.c_int32),
                ("device_visible_requests", C.c_int32), ("warmup_autotune", C.c_int32)]


class Request(C.Structure):
    _fields_ = [("request_id", C.c_int32), ("batch", C.c_int64), ("inputs", C.c_void_p * MAX_IO),
                ("outputs", C.c_void_p * MAX_IO), ("seq_len", C.c_int64)]


class Result(C.Structure):
    _fields_ = [("request_id", C.c_int32), ("ok", C.c_int32), ("rows", C.c_int64), ("batch_rows", C.c_int64),
                ("worker_id", C.c_int32), ("device_id", C.c_int32), ("executed_on", C.c_int32),
                ("codelet_ms", C.c_double), ("task_ms", C.c_double), ("latency_ms", C.c_double),
                ("first_value", C.c_float), ("error", C.c_char * 192), ("batch_seq", C.c_int64)]


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    capi.lib()  # libspu_b200.so first (libspu_host.so links against it via $ORIGIN)
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run `python -m starpu_inference_server_b200.build`")
    L = C.CDLL(LIB_PATH)
    vp, i32, i64, u64, u32, f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_uint32, C.c_float
    L.spu_server_create.argtypes = [C.POINTER(ServerCfg), C.POINTER(vp)]
    L.spu_server_create.restype = i32
    L.spu_server_destroy.argtypes = [vp]
    L.spu_server_destroy.restype = None
    L.spu_server_submit_batch.argtypes = [vp, C.POINTER(Request), i32, i32]
    L.spu_server_submit_batch.restype = i64
    L.spu_server_enqueue.argtypes = [vp, C.POINTER(Request)]
    L.spu_server_enqueue.restype = i64
    L.spu_server_wait.argtypes = [vp, i64, C.POINTER(Result), i32]
    L.spu_server_wait.restype = i32
    L.spu_server_wait_all.argtypes = [vp]
    L.spu_server_wait_all.restype = i32
    L.spu_server_stats.argtypes = [vp, C.POINTER(i64)]
    L.spu_server_stats.restype = None
    L.spu_host_last_error.restype = C.c_char_p
    L.spu_host_codelet_fields.argtypes = [C.POINTER(i64)]
    L.spu_host_codelet_fields.restype = None
    L.spu_host_buffer_byte_size.argtypes = [i32, u32, u64, i32, C.POINTER(u64)]
    L.spu_host_buffer_byte_size.restype = i32
    L.spu_host_select_replica.argtypes = [C.POINTER(i32), i32, C.POINTER(i32), i32, i32, i32, i32]
    L.spu_host_select_replica.restype = i32
    f64 = C.c_double
    L.spu_server_offered_load.argtypes = [vp, C.POINTER(Request), i32, i64, f64, u64, C.POINTER(f64), C.POINTER(f64),
                                          C.POINTER(f64), C.POINTER(i32), C.POINTER(f64)]
    L.spu_server_offered_load.restype = i32
    L.spu_host_request_table.argtypes = [C.POINTER(u64), C.POINTER(u64), C.POINTER(u64), C.POINTER(i64), C.POINTER(i64),
                                         i32, i64, u64, vp, u64, C.POINTER(i64)]
    L.spu_host_request_table.restype = i32
    L.spu_host_masks_all_ones.argtypes = [C.POINTER(vp), C.POINTER(i64), i32]
    L.spu_host_masks_all_ones.restype = i32
    L.spu_host_tile_aware_wants_more.argtypes = [i64, i64, i64, i32]
    L.spu_host_tile_aware_wants_more.restype = i32
    L.spu_host_nvtx_capture.argtypes = [i32, C.c_char_p, u64, C.POINTER(i64)]
    L.spu_host_nvtx_capture.restype = i32
    L.spu_host_replica_assignments.argtypes = [C.POINTER(i32), i32, i32, C.POINTER(i32), i32, C.POINTER(i32),
                                               C.POINTER(i32), i32, i32, i32, C.POINTER(i32)]
    L.spu_host_replica_assignments.restype = i32
    L.spu_host_gather_to_slot.argtypes = [C.POINTER(vp), C.POINTER(u64), i32, u64, vp, u64, C.POINTER(u64),
                                          C.POINTER(u64)]
    L.spu_host_gather_to_slot.restype = i32
    L.spu_host_slice_rows.argtypes = [u64, u64, i64, i64, C.POINTER(u64), C.POINTER(u64)]
    L.spu_host_slice_rows.restype = i64
    L.spu_host_can_merge.argtypes = [C.POINTER(i64), i32, i32, C.POINTER(i64), i32, i32]
    L.spu_host_can_merge.restype = i32
    L.spu_host_can_merge_varlen.argtypes = [C.POINTER(i64), i32, i32, C.POINTER(i64), i32, i32]
    L.spu_host_can_merge_varlen.restype = i32
    L.spu_host_gather_padded.argtypes = [C.POINTER(vp), C.POINTER(i64), i32, i64, vp, u64, C.POINTER(u64)]
    L.spu_host_gather_padded.restype = i32
    L.spu_host_call_cuda_func_affine.argtypes = [vp, vp, u64, f32, f32, i32, i32, vp, i32, i32, C.POINTER(i64)]
    L.spu_host_call_cuda_func_affine.restype = i32
    _lib = L
    return L


class HostError(RuntimeError):
    pass


def last_error() -> str:
    return lib().spu_host_last_error().decode(errors="replace")


def _tensor_cfg(name, dtype, dims):
    t = TensorCfg()
    t.name = name.encode()[:63]
    t.dtype = DT[dtype] if isinstance(dtype, str) else dtype
    t.ndims = len(dims)
    for i, d in enumerate(dims):
        t.dims[i] = d
    return t


class Server:
    """StarPU runtime + codelet + slot pools + batch collector around per-GPU replicas.

    ``inputs`` / ``outputs``: list of (name, dtype, dims) as models/<name>.yml declares them,
    dims[0] = per-request batch.  ``replicas``: list of replica handles (``BertReplica`` /
    ``VitReplica`` objects or raw pointers), one per entry of ``device_ids`` -- or, with
    ``per_worker_replicas`` (gpu_model_replication: per_worker), one per CUDA worker, device by device.
    """

    def __init__(self, inputs, outputs, device_ids=(0,), replicas=(), model_kind=MODEL_BERT, max_batch_size=32,
                 pool_size=12, nworker_per_cuda=1, cuda_pipeline=2, batch_coalesce_timeout_ms=0, n_workspaces=1,
                 affine=(1.0, 0.0), variable_length=False, length_bucket=0, per_worker_replicas=False, tile_rows=0,
                 device_visible_requests=False, warmup_autotune=False):
        cfg = ServerCfg()
        cfg.warmup_autotune = 1 if warmup_autotune else 0
        cfg.device_visible_requests = 1 if device_visible_requests else 0
        cfg.tile_rows = int(tile_rows)
        cfg.gpu_model_replication = 1 if per_worker_replicas else 0
        cfg.variable_length = 1 if variable_length else 0
        cfg.length_bucket = int(length_bucket)
        self.variable_length = bool(variable_length)
        cfg.n_devices = len(device_ids)
        for i, d in enumerate(device_ids):
            cfg.device_ids[i] = d
        cfg.nworker_per_cuda, cfg.cuda_pipeline = nworker_per_cuda, cuda_pipeline
        cfg.pool_size, cfg.max_batch_size = pool_size, max_batch_size
        cfg.batch_coalesce_timeout_ms = batch_coalesce_timeout_ms
        cfg.n_inputs, cfg.n_outputs = len(inputs), len(outputs)
        for i, t in enumerate(inputs):
            cfg.inputs[i] = _tensor_cfg(*t)
        for i, t in enumerate(outputs):
            cfg.outputs[i] = _tensor_cfg(*t)
        cfg.model_kind = model_kind
        self._keep = list(replicas)
        for i, r in enumerate(replicas):
            h = getattr(r, "_h", r)
            cfg.replicas[i] = h.value if isinstance(h, C.c_void_p) else h
        cfg.n_workspaces = n_workspaces
        cfg.affine_scale, cfg.affine_shift = affine
        self.inputs, self.outputs = list(inputs), list(outputs)
        h = C.c_void_p()
        if lib().spu_server_create(C.byref(cfg), C.byref(h)) != 0:
            raise HostError(last_error())
        self._h = h

    def _request(self, request_id, batch, in_arrays, out_arrays):
        rq = Request()
        rq.request_id, rq.batch = request_id, batch
        # variable-length servers: the request's own sequence length is dims[1] of its first input array
        rq.seq_len = int(in_arrays[0].shape[1]) if self.variable_length and len(in_arrays[0].shape) >= 2 else 0
        for i, a in enumerate(in_arrays):
            rq.inputs[i] = a.ctypes.data if hasattr(a, "ctypes") else a.data_ptr()
        for i, a in enumerate(out_arrays):
            rq.outputs[i] = None if a is None else (a.ctypes.data if hasattr(a, "ctypes") else a.data_ptr())
        return rq

    def submit_batch(self, requests, fixed_worker=-1):
        """requests: list of (request_id, batch, [input arrays], [output arrays]).  One StarPU task."""
        arr = (Request * len(requests))(*[self._request(*r) for r in requests])
        t = lib().spu_server_submit_batch(self._h, arr, len(requests), fixed_worker)
        if t < 0:
            raise HostError(last_error())
        return t

    def prepare_batch(self, requests):
        """Build the request table once (pointer table reuse across steps)."""
        return (Request * len(requests))(*[self._request(*r) for r in requests]), len(requests)

    def submit_prepared(self, prepared, fixed_worker=-1):
        arr, n = prepared
        t = lib().spu_server_submit_batch(self._h, arr, n, fixed_worker)
        if t < 0:
            raise HostError(last_error())
        return t

    def enqueue(self, request_id, batch, in_arrays, out_arrays):
        rq = self._request(request_id, batch, in_arrays, out_arrays)
        t = lib().spu_server_enqueue(self._h, C.byref(rq))
        if t < 0:
            raise HostError(last_error())
        return t

    def offered_load(self, prepared, n_req, rate_per_s, seed=1):
        """Open-loop Poisson arrivals of single requests through the batch collector (spu_server_offered_load).
        `prepared` = prepare_batch(...) rows to cycle over.  Returns dict of numpy arrays + elapsed seconds."""
        arr, n = prepared
        cod = np.zeros(n_req, dtype=np.float64)
        task = np.zeros(n_req, dtype=np.float64)
        lat = np.zeros(n_req, dtype=np.float64)
        rows = np.zeros(n_req, dtype=np.int32)
        el = C.c_double()
        f64p, i32p = C.POINTER(C.c_double), C.POINTER(C.c_int32)
        failed = lib().spu_server_offered_load(self._h, arr, n, n_req, float(rate_per_s), seed,
                                               cod.ctypes.data_as(f64p), task.ctypes.data_as(f64p), lat.ctypes.data_as(f64p),
                                               rows.ctypes.data_as(i32p), C.byref(el))
        if failed < 0:
            raise HostError(last_error())
        return {"failed": failed, "codelet_ms": cod, "task_ms": task, "latency_ms": lat, "batch_rows": rows, "elapsed_s": el.value}

    def wait(self, ticket, cap=64):
        res = (Result * cap)()
        n = lib().spu_server_wait(self._h, ticket, res, cap)
        if n < 0:
            raise HostError(last_error())
        return [res[i] for i in range(min(n, cap))]

    def wait_all(self):
        lib().spu_server_wait_all(self._h)

    def stats(self):
        out = (C.c_int64 * 6)()
        lib().spu_server_stats(self._h, out)
        return dict(zip(["tasks", "completed", "free_in", "free_out", "workers", "codelet_failures"], list(out)))

    def close(self):
        if getattr(self, "_h", None):
            lib().spu_server_destroy(self._h)
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


BERT_IO = dict(
    inputs=[("input_ids", "i64", [1, 128]), ("attention_mask", "i64", [1, 128])],
    outputs=[("output0", "f32", [1, 128, 768])],
)  # models/bert.yml:11-15
VIT_L_IO = dict(
    inputs=[("input", "f32", [1, 3, 224, 224])],
    outputs=[("output", "f32", [1, 1000])],
)  # models/vit_l_16.yml:13-16
