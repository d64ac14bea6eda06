"""CPU: host-side contract of the codelet wrapper (libspu_host.so) checked against the
reference's own known answers and the C oracle.  No CUDA call is made here.

Reference tests mirrored:
  codelet field values           tests/unit/core/unit_starpu_setup.cpp:4065-4074
  buffer_byte_size               src/core/starpu_setup.cpp:515-542 (+ unit tests around :4180)
  select_gpu_module              tests/unit/core/unit_starpu_setup.cpp:4106-4183
  gather bytes + nx              tests/unit/starpu/unit_starpu_task_runner_submission_and_slots.cpp:607-666
  scatter slices                 tests/unit/starpu/unit_starpu_task_runner_batching_and_internal.cpp:1514-1606
  can_merge_jobs                 src/starpu_task_worker/batch_composition_policy.cpp:102-151
"""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def host(built_libs):
    from starpu_inference_server_b200 import hostapi

    hostapi.lib()
    return hostapi


def test_host_header_symbols_all_exported(host):
    src = open(os.path.join(ROOT, "include", "spu_host.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    declared = sorted(set(re.findall(r"\b(spu_(?:server|host)_[a-z0-9_]+)\s*\(", src)))
    assert sorted(host.SYMBOLS) == declared
    for name in declared:
        assert hasattr(host.lib(), name), name


def test_codelet_fields_are_initialized(host):
    out = (C.c_int64 * 7)()
    host.lib().spu_host_codelet_fields(out)
    nbuffers, ctype, max_par, has_cpu, has_cuda, cuda_flags, where = list(out)
    assert nbuffers == -1          # STARPU_VARIABLE_NBUFFERS
    assert ctype == 2              # STARPU_FORKJOIN
    assert max_par == 2**31 - 1    # INT_MAX
    assert has_cuda == 1
    assert cuda_flags == 1         # STARPU_CUDA_ASYNC
    # deliberate difference from the reference (EXPECT_NE(cpu_funcs[0], nullptr)): the new
    # codelet has no CPU implementation to fall back to
    assert has_cpu == 0
    assert where == (1 << 3)       # STARPU_CUDA


@pytest.mark.parametrize("iface,nx,es", [(2, 4, 8), (2, 0, 4), (2, 4096 * 768, 4), (5, 123, 12), (5, 0, 0)])
def test_buffer_byte_size_matches_oracle(host, oracle_gs, iface, nx, es):
    got, want = C.c_uint64(), C.c_uint64()
    assert host.lib().spu_host_buffer_byte_size(iface, nx, es, 0, C.byref(got)) == 0
    oracle_gs.oracle_buffer_byte_size.argtypes = [C.c_int, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64)]
    assert oracle_gs.oracle_buffer_byte_size(iface, nx, es, C.byref(want)) == 0
    assert got.value == want.value


def test_buffer_byte_size_errors(host):
    out = C.c_uint64()
    assert host.lib().spu_host_buffer_byte_size(2, 1, 1, 1, C.byref(out)) == 1
    assert "StarPU buffer is null" in host.last_error()
    assert host.lib().spu_host_buffer_byte_size(0, 1, 1, 0, C.byref(out)) == 1  # matrix interface
    assert "Unsupported StarPU buffer interface id 0" in host.last_error()


def _select(host, device_ids, worker_ids, n_models, worker_id, device_id):
    d = (C.c_int32 * max(1, len(device_ids)))(*device_ids)
    w = (C.c_int32 * max(1, len(worker_ids)))(*worker_ids)
    return host.lib().spu_host_select_replica(d, len(device_ids), w, len(worker_ids), n_models, worker_id, device_id)


def test_select_gpu_module_lookup_order(host):
    # device id used as index when no table is given
    assert _select(host, [], [], 3, 0, 2) == 2
    # device_ids table: replica i serves device_ids[i]
    assert _select(host, [4, 7], [], 2, 0, 7) == 1
    assert _select(host, [4, 7], [], 2, 0, 4) == 0
    # per-worker table wins over the device table
    assert _select(host, [0], [10, 11, 12], 3, 11, 0) == 1


def test_select_gpu_module_throws_when_replica_missing(host):
    assert _select(host, [], [], 1, 0, 3) == -1
    assert "No GPU model replica available for device 3" in host.last_error()
    assert _select(host, [4], [], 1, 0, 5) == -1
    assert _select(host, [], [10, 11], 2, 12, 0) == -1
    assert "No GPU model replica available for worker 12 on device 0" in host.last_error()
    assert _select(host, [], [], 2, 0, -1) == -1


def _assign(host, device_ids, per_worker, worker_dev, pick_worker=-1, pick_device=-1):
    d = (C.c_int32 * max(1, len(device_ids)))(*device_ids)
    wd = (C.c_int32 * max(1, len(worker_dev)))(*worker_dev)
    od, ow, picked = (C.c_int32 * 32)(), (C.c_int32 * 32)(), C.c_int32(-2)
    n = host.lib().spu_host_replica_assignments(d, len(device_ids), 1 if per_worker else 0, wd, len(worker_dev), od, ow,
                                                32, pick_worker, pick_device, C.byref(picked))
    return n, [(od[i], ow[i]) for i in range(max(n, 0))], picked.value


def test_replica_registry_assignments_and_lookup(host):
    """build_gpu_replica_assignments (inference_runner.cpp:128-183) -> ReplicaRegistry -> fill_model_pointers
    (inference_task.cpp:469-516) -> select_gpu_module (starpu_setup.cpp:725-778)."""
    # per_device: one replica per non-negative device id, in configuration order; the codelet finds it by device
    n, a, picked = _assign(host, [2, 0, -1, 5], False, [], pick_device=5)
    assert (n, a, picked) == (3, [(2, -1), (0, -1), (5, -1)], 2)
    # per_worker: one per CUDA worker of every listed device, devices in order, workers in id order
    worker_dev = [0, 0, 1, 1, 3]  # StarPU worker w runs on device worker_dev[w]
    n, a, picked = _assign(host, [1, 0], True, worker_dev, pick_worker=1, pick_device=0)
    assert (n, a) == (4, [(1, 2), (1, 3), (0, 0), (0, 1)]) and picked == 3
    # a worker that has no replica is an error of the codelet, as in the reference
    n, a, picked = _assign(host, [1], True, worker_dev, pick_worker=0, pick_device=0)
    assert n == -1 and "No GPU model replica available for worker 0 on device 0" in host.last_error()
    # a listed device without a CUDA worker cannot be served per worker
    n, _, _ = _assign(host, [0, 2], True, worker_dev)
    assert n == -1 and "No CUDA workers available for device 2" in host.last_error()
    assert _assign(host, [], False, [])[0] == 0


def _gather(host, arrays, elemsize, capacity):
    n = len(arrays)
    srcs = (C.c_void_p * max(1, n))(*[a.ctypes.data for a in arrays])
    nbytes = (C.c_uint64 * max(1, n))(*[a.nbytes for a in arrays])
    slot = np.full(capacity, 0xEE, np.uint8)
    nx, total = C.c_uint64(), C.c_uint64()
    rc = host.lib().spu_host_gather_to_slot(srcs, nbytes, n, elemsize, slot.ctypes.data, capacity, C.byref(nx),
                                            C.byref(total))
    return rc, slot, nx.value, total.value


def test_gather_known_answer(host):
    """jobs {1,2} + {3,4} -> slot 1,2,3,4 and nx == 4 (reference test :607-666)."""
    a, b = np.array([1, 2], np.float32), np.array([3, 4], np.float32)
    rc, slot, nx, total = _gather(host, [a, b], 4, 64)
    assert rc == 0 and nx == 4 and total == 16
    assert slot[:16].view(np.float32).tolist() == [1, 2, 3, 4]
    assert (slot[16:] == 0xEE).all()


def test_gather_matches_oracle_bytes(host, oracle_gs):
    rng = np.random.default_rng(5)
    arrays = [rng.integers(0, 30522, size=(b, 128), dtype=np.int64) for b in (1, 3, 1, 2, 1)]
    cap = 32 * 128 * 8
    rc, slot, nx, total = _gather(host, arrays, 8, cap)
    assert rc == 0
    n = len(arrays)
    srcs = (C.c_void_p * n)(*[a.ctypes.data for a in arrays])
    nbytes = (C.c_size_t * n)(*[a.nbytes for a in arrays])
    numel = (C.c_size_t * n)(*[a.size for a in arrays])
    want = np.full(cap, 0xEE, np.uint8)
    onx, obytes = C.c_size_t(), C.c_size_t()
    oracle_gs.oracle_gather_concat.argtypes = [C.POINTER(C.c_void_p), C.POINTER(C.c_size_t), C.POINTER(C.c_size_t),
                                               C.c_size_t, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t),
                                               C.POINTER(C.c_size_t)]
    assert oracle_gs.oracle_gather_concat(srcs, nbytes, numel, n, want.ctypes.data, cap, C.byref(onx),
                                          C.byref(obytes)) == 0
    assert (slot == want).all() and nx == onx.value == 8 * 128 and total == obytes.value


def _gather_padded(host, rows, seq, cap_rows):
    n = len(rows)
    srcs = (C.c_void_p * n)(*[r.ctypes.data for r in rows])
    lens = (C.c_int64 * n)(*[r.size for r in rows])
    slot = np.full(cap_rows * seq, -77, np.int64)
    nx = C.c_uint64()
    rc = host.lib().spu_host_gather_padded(srcs, lens, n, seq, slot.ctypes.data, slot.nbytes, C.byref(nx))
    return rc, slot, nx.value


def test_varlen_gather_known_answer(host):
    """requests {1,2,3} + {4} padded to the batch max 3 -> 1,2,3,4,0,0 and nx == 6."""
    rc, slot, nx = _gather_padded(host, [np.array([1, 2, 3], np.int64), np.array([4], np.int64)], 3, 4)
    assert rc == 0 and nx == 6
    assert slot[:6].tolist() == [1, 2, 3, 4, 0, 0] and (slot[6:] == -77).all()


def test_varlen_gather_matches_oracle(host, oracle_gs):
    """The padded host gather of ids and of an all-ones mask equals the C oracle's gather+pad+mask."""
    rng = np.random.default_rng(9)
    lens = [128, 1, 77, 16, 100]
    seq = max(lens)
    ids = [rng.integers(1, 30522, size=n, dtype=np.int64) for n in lens]
    ones = [np.ones(n, np.int64) for n in lens]
    rc, got_ids, nx = _gather_padded(host, ids, seq, 8)
    rc2, got_mask, _ = _gather_padded(host, ones, seq, 8)
    assert rc == 0 and rc2 == 0 and nx == len(lens) * seq
    n = len(lens)
    PP = C.c_void_p * n
    want_ids, want_mask = np.zeros((n, seq), np.int64), np.zeros((n, seq), np.int64)
    lens32 = np.array(lens, np.int32)
    # no argtypes on the shared oracle handle (other tests call it with explicit ctypes values)
    assert oracle_gs.oracle_gather_pad_mask(PP(*[a.ctypes.data for a in ids]), None, lens32.ctypes.data_as(C.c_void_p),
                                            C.c_size_t(n), C.c_size_t(seq), C.c_int64(0),
                                            want_ids.ctypes.data_as(C.c_void_p),
                                            want_mask.ctypes.data_as(C.c_void_p)) == 0
    assert (got_ids[:n * seq].reshape(n, seq) == want_ids).all()
    assert (got_mask[:n * seq].reshape(n, seq) == want_mask).all()


def test_varlen_gather_rejects_long_and_overflow(host):
    rc, _, _ = _gather_padded(host, [np.arange(5, dtype=np.int64)], 4, 4)   # request longer than the batch seq
    assert rc == 1 and "variable-length input" in host.last_error()
    rc, _, _ = _gather_padded(host, [np.arange(4, dtype=np.int64)] * 3, 4, 2)  # 3 rows into a 2-row slot
    assert rc == 1 and "exceeds allocated slot buffer capacity" in host.last_error()


def test_can_merge_jobs_variable_length(host):
    I64 = 3
    m = lambda da, db, tb=I64: host.lib().spu_host_can_merge_varlen(  # noqa: E731
        (C.c_int64 * len(da))(*da), len(da), I64, (C.c_int64 * len(db))(*db), len(db), tb)
    assert m([1, 128], [1, 64]) == 1          # only the sequence dim is relaxed
    assert m([1, 128], [3, 17]) == 1
    assert m([1, 128, 4], [1, 64, 4]) == 1
    assert m([1, 128, 4], [1, 64, 5]) == 0    # trailing dims past the sequence still have to match
    assert m([1, 128], [1, 64], 0) == 0       # dtype
    assert m([1, 128], [1, 64, 1]) == 0       # rank


def test_gather_capacity_exception(host):
    a = np.zeros(8, np.float32)
    rc, _, _, _ = _gather(host, [a, a, a], 4, 64)
    assert rc == 1 and "exceeds allocated slot buffer capacity" in host.last_error()


@pytest.mark.parametrize("rows,off,bs", [(3, 0, 1), (3, 1, 2), (3, 2, 5), (3, 3, 1), (32, 7, 0), (1, 0, 1), (0, 0, 1)])
def test_slice_rows_matches_oracle(host, oracle_gs, rows, off, bs):
    row_bytes = 4000
    o, ln = C.c_uint64(), C.c_uint64()
    processed = host.lib().spu_host_slice_rows(rows, row_bytes, off, bs, C.byref(o), C.byref(ln))
    agg = np.arange(max(rows, 1) * row_bytes, dtype=np.uint8)
    dst = np.zeros(max(rows, 1) * row_bytes + 8, np.uint8)
    got_rows = C.c_int64()
    oracle_gs.oracle_scatter_slice.argtypes = [C.c_void_p, C.c_size_t, C.c_size_t, C.c_int64, C.c_int64, C.c_void_p,
                                               C.POINTER(C.c_int64)]
    assert oracle_gs.oracle_scatter_slice(agg.ctypes.data, rows, rows * row_bytes, off, bs, dst.ctypes.data,
                                          C.byref(got_rows)) == 0
    assert ln.value == got_rows.value * row_bytes
    if got_rows.value > 0:
        assert processed == got_rows.value and o.value == off * row_bytes
        assert (dst[:ln.value] == agg[o.value:o.value + ln.value]).all()
    else:
        assert processed == max(1, bs)  # reference: processed_length stays slice_size when nothing is cut


def test_slice_known_answer(host):
    """batch sizes 1 + 2 over 3 rows -> rows [0,1) and [1,3) (reference test :1514-1606)."""
    o, ln = C.c_uint64(), C.c_uint64()
    assert host.lib().spu_host_slice_rows(3, 8, 0, 1, C.byref(o), C.byref(ln)) == 1 and (o.value, ln.value) == (0, 8)
    assert host.lib().spu_host_slice_rows(3, 8, 1, 2, C.byref(o), C.byref(ln)) == 2 and (o.value, ln.value) == (8, 16)


def _merge(host, da, ta, db, tb):
    a, b = (C.c_int64 * len(da))(*da), (C.c_int64 * len(db))(*db)
    return host.lib().spu_host_can_merge(a, len(da), ta, b, len(db), tb)


def test_can_merge_jobs(host):
    I64, F32 = 3, 0
    assert _merge(host, [1, 128], I64, [1, 128], I64) == 1
    assert _merge(host, [1, 128], I64, [4, 128], I64) == 1    # batch dim is free
    assert _merge(host, [1, 128], I64, [1, 64], I64) == 0     # trailing dims must match
    assert _merge(host, [1, 128], I64, [1, 128], F32) == 0    # dtype must match
    assert _merge(host, [1, 3, 224, 224], F32, [2, 3, 224, 224], F32) == 1
    assert _merge(host, [1, 128], I64, [1, 128, 1], I64) == 0  # rank must match
    assert _merge(host, [5], F32, [9], F32) == 1              # rank 1: nothing but dim 0


def test_tile_aware_close_rule(host):
    """The collector's tile-aware preference (spu_server_cfg.tile_rows): a batch that ends inside a 256-row GEMM tile
    and may still grow asks for more; whole tiles, a full batch, or tile_rows = 0 do not."""
    L = host.lib()
    assert L.spu_host_tile_aware_wants_more(3, 128, 32, 256) == 1     # 384 token rows = 1.5 tiles
    assert L.spu_host_tile_aware_wants_more(4, 128, 32, 256) == 0     # 512 = 2 tiles
    assert L.spu_host_tile_aware_wants_more(31, 128, 32, 256) == 1
    assert L.spu_host_tile_aware_wants_more(32, 128, 32, 256) == 0    # at the cap
    assert L.spu_host_tile_aware_wants_more(5, 100, 64, 256) == 1     # 500 rows
    assert L.spu_host_tile_aware_wants_more(64, 100, 64, 256) == 0    # at the cap although 6400 % 256 != 0
    assert L.spu_host_tile_aware_wants_more(3, 128, 32, 0) == 0       # switched off
    assert L.spu_host_tile_aware_wants_more(1, 197, 64, 256) == 1     # ViT: 197 tokens per image


def test_request_table_layout_for_device_visible_requests(host):
    """write_request_table: one entry per SAMPLE ROW in request order -- ids / mask addresses stride
    by the request's own length, the destination by length x token bytes -- int32 lengths behind the three pointer
    arrays, the table handle resized to the table, the second input to one element; ragged batches scatter live rows."""
    import ctypes as C

    L = host.lib()
    u64, i64 = C.c_uint64, C.c_int64
    batch, lens = [1, 2, 1], [5, 9, 9]
    ids, mask, dst = [0x1000, 0x2000, 0x3000], [0x11000, 0x12000, 0x13000], [0x100000, 0x200000, 0x300000]
    slot = np.zeros(64, dtype=np.int64)
    out = (i64 * 8)()
    tb = 768 * 4
    rc = L.spu_host_request_table((u64 * 3)(*ids), (u64 * 3)(*mask), (u64 * 3)(*dst), (i64 * 3)(*batch), (i64 * 3)(*lens),
                                  3, 9, tb, slot.ctypes.data, slot.nbytes, out)
    assert rc == 0, host.last_error()
    rows, off_mask, off_dst, off_len, table_bytes, full_rows, nx0, nx1 = list(out)
    assert (rows, off_mask, off_dst, off_len) == (4, 32, 64, 96) and table_bytes == 112 and full_rows == 0
    assert nx0 == table_bytes // 8 and nx1 == 1
    raw = slot.view(np.uint8)
    tab = lambda off: raw[off:off + 32].view(np.uint64).tolist()
    assert tab(0) == [0x1000, 0x2000, 0x2000 + 9 * 8, 0x3000]
    assert tab(off_mask) == [0x11000, 0x12000, 0x12000 + 9 * 8, 0x13000]
    assert tab(off_dst) == [0x100000, 0x200000, 0x200000 + 9 * tb, 0x300000]
    assert raw[off_len:off_len + 16].view(np.int32).tolist() == [5, 9, 9, 9]
    # equal lengths = the batch's seq: full-row scatter
    rc = L.spu_host_request_table((u64 * 2)(*ids[:2]), (u64 * 2)(*mask[:2]), (u64 * 2)(*dst[:2]), (i64 * 2)(1, 1),
                                  (i64 * 2)(9, 9), 2, 9, tb, slot.ctypes.data, slot.nbytes, out)
    assert rc == 0 and out[5] == 1
    # a request without a response buffer, a request longer than the batch, a slot too small for the table
    rc = L.spu_host_request_table((u64 * 1)(ids[0]), (u64 * 1)(mask[0]), (u64 * 1)(0), (i64 * 1)(1), (i64 * 1)(5), 1, 9, tb,
                                  slot.ctypes.data, slot.nbytes, out)
    assert rc == 1 and "output buffer" in host.last_error()
    rc = L.spu_host_request_table((u64 * 1)(ids[0]), (u64 * 1)(mask[0]), (u64 * 1)(dst[0]), (i64 * 1)(1), (i64 * 1)(10), 1, 9,
                                  tb, slot.ctypes.data, slot.nbytes, out)
    assert rc == 1 and "longer than the batch" in host.last_error()
    rc = L.spu_host_request_table((u64 * 3)(*ids), (u64 * 3)(*mask), (u64 * 3)(*dst), (i64 * 3)(*batch), (i64 * 3)(*lens),
                                  3, 9, tb, slot.ctypes.data, 64, out)
    assert rc == 1 and "capacity" in host.last_error()


def test_masks_all_ones_decides_the_packed_forward(host):
    import ctypes as C

    L = host.lib()
    a, b = np.ones(7, dtype=np.int64), np.ones(12, dtype=np.int64)
    ptrs = (C.c_void_p * 2)(a.ctypes.data, b.ctypes.data)
    n = (C.c_int64 * 2)(7, 12)
    assert L.spu_host_masks_all_ones(ptrs, n, 2) == 1
    b[11] = 0  # a client-side padded mask: lengths are not the requests' dims[1] any more
    assert L.spu_host_masks_all_ones(ptrs, n, 2) == 0
    b[11], a[0] = 3, -1  # any non-zero word keeps a token
    assert L.spu_host_masks_all_ones(ptrs, n, 2) == 1
