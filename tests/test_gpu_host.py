"""GPU: the codelet wrapper and the task path around it (libspu_host.so), through the same
entry a StarPU CUDA worker uses: codelet.cuda_funcs[0](buffers, cl_arg).

Reference tests mirrored:
  cuda_func x+1.5 on device buffers + side effects   tests/unit/core/unit_starpu_setup.cpp:2332-2433
  missing replica -> StarPUCodeletException          tests/unit/core/unit_starpu_setup.cpp:4106-4183
  cpu/cuda_func x+1 on {1,2,3} -> {2,3,4}            tests/integration/starpu/integration_starpu_setup.cpp:42-60
  x*2 through the runner                             tests/integration/core/integration_inference_runner.cpp:72-120
  batched sub-jobs sliced back in order              tests/unit/starpu/unit_starpu_task_runner_batching_and_internal.cpp:1514-1606
"""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def host(built_libs):
    from starpu_inference_server_b200 import capi, hostapi

    assert torch.cuda.is_available() and capi.lib().spu_device_is_sm100(0) == 1
    hostapi.lib()
    return hostapi


@pytest.mark.parametrize("vector_iface", [0, 1])
def test_cuda_func_copies_results_to_device_buffer(host, vector_iface):
    x = torch.tensor([1.0, 2.0, 3.0], device="cuda")
    y = torch.zeros(3, device="cuda")
    s = torch.cuda.Stream()
    rec = (C.c_int64 * 6)()
    label, counts = C.create_string_buffer(96), (C.c_int64 * 2)()
    host.lib().spu_host_nvtx_capture(1, label, 96, counts)
    before = list(counts)
    rc = host.lib().spu_host_call_cuda_func_affine(x.data_ptr(), y.data_ptr(), 3, 1.0, 1.5, 0, 77, s.cuda_stream, 1,
                                                   vector_iface, rec)
    host.lib().spu_host_nvtx_capture(0, label, 96, counts)  # back to the real NVTX calls
    assert rc == 0, host.last_error()
    # one NVTX range around the codelet body, named like the reference's (starpu_setup.cpp:817)
    assert label.value == b"codelet.cuda job 0 dev 0" and [counts[0] - before[0], counts[1] - before[1]] == [1, 1]
    s.synchronize()
    assert y.cpu().tolist() == [2.5, 3.5, 4.5]
    device_id, worker_id, executed_on, start_set, end_set, gauge = list(rec)
    assert (device_id, worker_id, executed_on) == (0, 77, 1)  # DeviceType::CUDA
    assert start_set == 1 and end_set == 1 and gauge == 0


def test_cuda_func_wraps_missing_replica(host):
    x = torch.ones(3, device="cuda")
    y = torch.zeros(3, device="cuda")
    rec = (C.c_int64 * 6)()
    rc = host.lib().spu_host_call_cuda_func_affine(x.data_ptr(), y.data_ptr(), 3, 1.0, 1.0, 0, 5,
                                                   torch.cuda.current_stream().cuda_stream, 0, 0, rec)
    assert rc == 1 and "No GPU model replica available for device 0" in host.last_error()
    torch.cuda.synchronize()
    assert y.cpu().tolist() == [0, 0, 0]


def _affine_server(host, scale, shift, **kw):
    return host.Server(inputs=[("x", "f32", [1, 3])], outputs=[("y", "f32", [1, 3])], model_kind=host.MODEL_AFFINE,
                       affine=(scale, shift), **kw)


def test_task_path_x_plus_one(host):
    with _affine_server(host, 1.0, 1.0, max_batch_size=4, pool_size=2) as srv:
        x = np.array([[1, 2, 3]], np.float32)
        y = np.zeros((1, 3), np.float32)
        (r,) = srv.wait(srv.submit_batch([(7, 1, [x], [y])]))
        assert r.ok == 1 and r.request_id == 7 and r.rows == 1 and r.batch_rows == 1
        assert y.tolist() == [[2, 3, 4]]
        assert r.executed_on == 1 and r.device_id == 0 and r.worker_id >= 0
        assert r.codelet_ms > 0 and r.task_ms > 0 and r.latency_ms >= r.task_ms
        st = srv.stats()
        assert st["tasks"] == 1 and st["completed"] == 1 and st["free_in"] == 2 and st["free_out"] == 2


def test_batched_sub_jobs_are_gathered_and_sliced_in_order(host):
    """master b=1 + sub-jobs b=2, b=1: rows are concatenated in request order and each
    request gets exactly its own rows back (x*2, the reference's integration toy)."""
    with _affine_server(host, 2.0, 0.0, max_batch_size=8, pool_size=3, nworker_per_cuda=2) as srv:
        rng = np.random.default_rng(0)
        xs = [rng.standard_normal((b, 3)).astype(np.float32) for b in (1, 2, 1)]
        ys = [np.zeros_like(x) for x in xs]
        res = srv.wait(srv.submit_batch([(10 + i, x.shape[0], [x], [y]) for i, (x, y) in enumerate(zip(xs, ys))]))
        assert [r.request_id for r in res] == [10, 11, 12]
        assert [r.rows for r in res] == [1, 2, 1] and all(r.batch_rows == 4 and r.ok == 1 for r in res)
        for x, y in zip(xs, ys):
            assert (y == x * 2).all()


def test_batch_over_capacity_is_rejected(host):
    with _affine_server(host, 1.0, 0.0, max_batch_size=2, pool_size=1) as srv:
        x = np.zeros((3, 3), np.float32)
        with pytest.raises(host.HostError, match="exceeds input pool capacity"):
            srv.submit_batch([(0, 3, [x], [np.zeros_like(x)])])
        assert srv.stats()["free_in"] == 1  # the slot went back


def test_dynamic_batching_and_many_workers(host):
    with _affine_server(host, 1.0, 0.5, max_batch_size=8, pool_size=4, nworker_per_cuda=4, cuda_pipeline=4,
                        batch_coalesce_timeout_ms=2) as srv:
        n = 200
        xs = [np.full((1, 3), float(i), np.float32) for i in range(n)]
        ys = [np.zeros((1, 3), np.float32) for _ in range(n)]
        tickets = [srv.enqueue(i, 1, [xs[i]], [ys[i]]) for i in range(n)]
        workers, sizes = set(), []
        for i, t in enumerate(tickets):
            (r,) = srv.wait(t)
            assert r.ok == 1 and r.request_id == i
            workers.add(r.worker_id)
            sizes.append(r.batch_rows)
            assert (ys[i] == i + 0.5).all()
        st = srv.stats()
        assert st["completed"] == n and st["tasks"] <= n and st["workers"] == 4
        assert max(sizes) > 1 and max(sizes) <= 8  # the collector did coalesce
        assert st["free_in"] == 4 and st["free_out"] == 4 and st["codelet_failures"] == 0


@pytest.fixture(scope="module")
def bert_tiny(host):
    from oracle import bert_oracle as bo
    from starpu_inference_server_b200.bert import BertReplica

    spec = bo.BERT_TINY
    w = bo.make_weights(spec, 0)
    rep = BertReplica(w, max_batch=8, max_seq=32, n_workspaces=2)
    yield bo, spec, w, rep
    rep.close()


def test_bert_through_the_task_path(host, bert_tiny):
    """gRPC-shaped requests ([1,S] ids+mask each) -> host gather -> StarPU task -> codelet ->
    C-ABI kernels -> D2H -> slices.  Must equal the direct C-ABI forward bit for bit and the
    fp32 oracle within the bf16 tolerance."""
    bo, spec, w, rep = bert_tiny
    S, H = 32, spec.hidden
    ids, mask = bo.make_inputs(spec, 5, S, seed=3, min_len=4)
    ref = bo.bert_forward(w, spec, ids, mask)
    direct = torch.empty(5, S, H, device="cuda")
    rep.forward_async(torch.from_numpy(ids).cuda(), torch.from_numpy(mask).cuda(), direct, ws=1)
    torch.cuda.synchronize()
    with host.Server(inputs=[("input_ids", "i64", [1, S]), ("attention_mask", "i64", [1, S])],
                     outputs=[("output0", "f32", [1, S, H])], replicas=[rep], model_kind=host.MODEL_BERT,
                     max_batch_size=8, pool_size=2, n_workspaces=2) as srv:
        outs = [np.zeros((1, S, H), np.float32) for _ in range(5)]
        reqs = [(i, 1, [ids[i:i + 1].copy(), mask[i:i + 1].copy()], [outs[i]]) for i in range(5)]
        res = srv.wait(srv.submit_batch(reqs))
        assert all(r.ok == 1 and r.batch_rows == 5 for r in res)
        got = np.concatenate(outs, 0)
        assert (got == direct.cpu().numpy()).all()
        assert np.abs(got - ref).max() / np.abs(ref).max() < 2e-2
        # wrong sequence length for this server's config: the codelet's size check fires and
        # the failure reaches the request as ok == 0 (gRPC INTERNAL in the reference)
    with host.Server(inputs=[("input_ids", "i64", [1, S]), ("attention_mask", "i64", [1, S])],
                     outputs=[("output0", "f32", [1, S, H + 8])], replicas=[rep], model_kind=host.MODEL_BERT,
                     max_batch_size=8, pool_size=2, n_workspaces=2) as srv:
        out = np.zeros((1, S, H + 8), np.float32)
        (r,) = srv.wait(srv.submit_batch([(0, 1, [ids[:1].copy(), mask[:1].copy()], [out])]))
        assert r.ok == 0 and srv.stats()["codelet_failures"] == 1


def test_variable_length_requests_through_the_task_path(host, bert_tiny):
    """SURVEY 8(f) row 1 on the served path: requests of different lengths merge into one task, the host gather pads
    them to the batch's longest request (mask synthesised by the zero padding), the codelet runs [B, S_batch], and
    every request gets exactly its own rows back.  Each response equals the direct C-ABI forward on the padded batch
    bit for bit, and the fp32 oracle run on THAT REQUEST ALONE (its own length, no padding) within the tolerance."""
    bo, spec, w, rep = bert_tiny
    S_max, H = 32, spec.hidden
    rng = np.random.default_rng(21)
    lens = [17, 32, 1, 9, 24]
    ids = [rng.integers(1, spec.vocab, size=(1, n), dtype=np.int64) for n in lens]
    ones = [np.ones((1, n), np.int64) for n in lens]
    S_b = max(lens)
    ids_pad, mask_pad = np.zeros((len(lens), S_b), np.int64), np.zeros((len(lens), S_b), np.int64)
    for i, n in enumerate(lens):
        ids_pad[i, :n], mask_pad[i, :n] = ids[i][0], 1
    direct = torch.empty(len(lens), S_b, H, device="cuda")
    rep.forward_async(torch.from_numpy(ids_pad).cuda(), torch.from_numpy(mask_pad).cuda(), direct, ws=1)
    torch.cuda.synchronize()
    direct = direct.cpu().numpy()
    with host.Server(inputs=[("input_ids", "i64", [1, S_max]), ("attention_mask", "i64", [1, S_max])],
                     outputs=[("output0", "f32", [1, S_max, H])], replicas=[rep], model_kind=host.MODEL_BERT,
                     max_batch_size=8, pool_size=2, n_workspaces=2, variable_length=True) as srv:
        outs = [np.full((1, n, H), np.nan, np.float32) for n in lens]
        reqs = [(i, 1, [ids[i], ones[i]], [outs[i]]) for i in range(len(lens))]
        res = srv.wait(srv.submit_batch(reqs))
        assert all(r.ok == 1 and r.batch_rows == len(lens) for r in res)
        for i, n in enumerate(lens):
            assert (outs[i][0] == direct[i, :n]).all(), f"request {i}"
            alone = bo.bert_forward(w, spec, ids[i], ones[i])
            assert np.abs(outs[i] - alone).max() / np.abs(alone).max() < 2e-2, f"request {i} vs oracle"
        # a shorter batch lands a smaller D2H: second task holds only short requests
        short = [np.full((1, n, H), np.nan, np.float32) for n in (3, 5)]
        res = srv.wait(srv.submit_batch([(10, 1, [ids[0][:, :3].copy(), ones[0][:, :3].copy()], [short[0]]),
                                         (11, 1, [ids[1][:, :5].copy(), ones[1][:, :5].copy()], [short[1]])]))
        assert all(r.ok == 1 for r in res) and np.isfinite(short[0]).all() and np.isfinite(short[1]).all()
        # dynamic batching merges different lengths too
        outs2 = [np.full((1, n, H), np.nan, np.float32) for n in lens]
        tickets = [srv.enqueue(100 + i, 1, [ids[i], ones[i]], [outs2[i]]) for i in range(len(lens))]
        for t in tickets:
            (r,) = srv.wait(t)
            assert r.ok == 1
        for i in range(len(lens)):
            alone = bo.bert_forward(w, spec, ids[i], ones[i])
            assert np.abs(outs2[i] - alone).max() / np.abs(alone).max() < 2e-2
        # longer than the configured cap: refused before anything is copied
        too_long = np.ones((1, S_max + 1), np.int64)
        with pytest.raises(host.HostError):
            srv.wait(srv.submit_batch([(0, 1, [too_long, too_long], [None])]))


def test_variable_length_multi_sample_request_is_compacted_to_its_own_shape(host, bert_tiny):
    """A request carrying 2 samples of length 9 merged with a length-30 request: its response buffer is sized by ITS
    shape [2, 9, H]; the copy must be compact (sample by sample), never the padded [2, 30, H] -- a sentinel right
    behind the buffer stays untouched (heap overflow in the response memcpy)."""
    bo, spec, w, rep = bert_tiny
    S_max, H = 32, spec.hidden
    rng = np.random.default_rng(8)
    ids2 = rng.integers(1, spec.vocab, size=(2, 9), dtype=np.int64)
    ids1 = rng.integers(1, spec.vocab, size=(1, 30), dtype=np.int64)
    with host.Server(inputs=[("input_ids", "i64", [1, S_max]), ("attention_mask", "i64", [1, S_max])],
                     outputs=[("output0", "f32", [1, S_max, H])], replicas=[rep], model_kind=host.MODEL_BERT,
                     max_batch_size=8, pool_size=2, n_workspaces=2, variable_length=True) as srv:
        backing = np.full(2 * 9 * H + 4096, -7.0, np.float32)
        out2 = backing[: 2 * 9 * H].reshape(2, 9, H)
        out1 = np.full((1, 30, H), np.nan, np.float32)
        res = srv.wait(srv.submit_batch([(0, 2, [ids2, np.ones_like(ids2)], [out2]),
                                         (1, 1, [ids1, np.ones_like(ids1)], [out1])]))
        assert all(r.ok == 1 for r in res) and res[0].batch_seq == 30 and res[0].rows == 2
        assert (backing[2 * 9 * H:] == -7.0).all(), "response copy ran past the request's own [2, 9, H] buffer"
        for k in range(2):
            alone = bo.bert_forward(w, spec, ids2[k:k + 1], np.ones((1, 9), np.int64))
            assert np.abs(out2[k:k + 1] - alone).max() / np.abs(alone).max() < 2e-2, k
        alone = bo.bert_forward(w, spec, ids1, np.ones_like(ids1))
        assert np.abs(out1 - alone).max() / np.abs(alone).max() < 2e-2


def test_bf16_response_through_the_task_path(host, bert_tiny):
    """models/<name>.yml output declared TYPE_BF16 + a replica created with SPU_OUTPUT_BF16: the slot, the D2H and
    the per-request slices carry 2 bytes per element; values = the fp32 path's rounded to bf16."""
    from starpu_inference_server_b200 import capi
    from starpu_inference_server_b200.bert import BertReplica

    bo, spec, w, rep32 = bert_tiny
    S, H = 32, spec.hidden
    ids, mask = bo.make_inputs(spec, 4, S, seed=5, min_len=4)
    direct = torch.empty(4, S, H, device="cuda")
    rep32.forward_async(torch.from_numpy(ids).cuda(), torch.from_numpy(mask).cuda(), direct, ws=1)
    torch.cuda.synchronize()
    want = direct.to(torch.bfloat16).cpu().view(torch.int16).numpy()
    rep16 = BertReplica(w, max_batch=8, max_seq=32, n_workspaces=2, flags=capi.SPU_OUTPUT_BF16)
    with host.Server(inputs=[("input_ids", "i64", [1, S]), ("attention_mask", "i64", [1, S])],
                     outputs=[("output0", "bf16", [1, S, H])], replicas=[rep16], model_kind=host.MODEL_BERT,
                     max_batch_size=8, pool_size=2, n_workspaces=2) as srv:
        outs = [np.zeros((1, S, H), np.int16) for _ in range(4)]
        res = srv.wait(srv.submit_batch([(i, 1, [ids[i:i + 1].copy(), mask[i:i + 1].copy()], [outs[i]])
                                         for i in range(4)]))
        assert all(r.ok == 1 for r in res)
        assert (np.concatenate(outs, 0) == want).all()
    rep16.close()


def test_length_bucketing_in_the_collector(host, bert_tiny):
    """variable_length + length_bucket: requests arriving short/long interleaved are batched per bucket of
    ceil(L / 16) (each task pads to its own bucket's longest request), every response still equals the fp32 oracle
    on that request alone, and the oldest request always opens the next batch."""
    bo, spec, w, rep = bert_tiny
    S_max, H = 32, spec.hidden
    rng = np.random.default_rng(4)
    lens = [5, 30, 6, 29, 7, 31, 4, 28]
    ids = [rng.integers(1, spec.vocab, size=(1, n), dtype=np.int64) for n in lens]
    ones = [np.ones((1, n), np.int64) for n in lens]
    with host.Server(inputs=[("input_ids", "i64", [1, S_max]), ("attention_mask", "i64", [1, S_max])],
                     outputs=[("output0", "f32", [1, S_max, H])], replicas=[rep], model_kind=host.MODEL_BERT,
                     max_batch_size=4, pool_size=2, n_workspaces=2, batch_coalesce_timeout_ms=200,
                     variable_length=True, length_bucket=16) as srv:
        outs = [np.full((1, n, H), np.nan, np.float32) for n in lens]
        tickets = [srv.enqueue(i, 1, [ids[i], ones[i]], [outs[i]]) for i in range(len(lens))]
        res = [srv.wait(t)[0] for t in tickets]
        assert all(r.ok == 1 for r in res)
        for i, (n, r) in enumerate(zip(lens, res)):
            assert r.batch_rows == 4, (i, r.batch_rows)
            assert r.batch_seq == (7 if n <= 16 else 31), (i, n, r.batch_seq)  # padded to its own bucket's longest
            alone = bo.bert_forward(w, spec, ids[i], ones[i])
            assert np.abs(outs[i] - alone).max() / np.abs(alone).max() < 2e-2
        assert srv.stats()["tasks"] == 2


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
def test_two_devices_in_one_process(host):
    """use_cuda: [{device_ids: [0, 1]}] (config_loader_device.hpp:104-149): one replica per device, one server; tasks
    land on both devices and every response equals the single-device result bit for bit (kernel attributes such as
    the dynamic shared-memory opt-in are per device)."""
    from oracle import bert_oracle as bo
    from starpu_inference_server_b200.bert import BertReplica

    spec = bo.BERT_TINY
    w = bo.make_weights(spec, 0)
    S, H = 160, spec.hidden  # > 128: the KV-loop attention variant (101 KB of dynamic smem) on both devices
    reps = [BertReplica(w, device=d, max_batch=4, max_seq=S, n_workspaces=1) for d in (0, 1)]
    ids, mask = bo.make_inputs(spec, 4, S, seed=13, min_len=20)
    direct = torch.empty(4, S, H, device="cuda:0")
    reps[0].forward_async(torch.from_numpy(ids).cuda(0), torch.from_numpy(mask).cuda(0), direct)
    torch.cuda.synchronize(0)
    want = direct.cpu().numpy()
    with host.Server(inputs=[("input_ids", "i64", [1, S]), ("attention_mask", "i64", [1, S])],
                     outputs=[("output0", "f32", [1, S, H])], replicas=reps, device_ids=[0, 1],
                     model_kind=host.MODEL_BERT, max_batch_size=4, pool_size=4, n_workspaces=1) as srv:
        devices = set()
        for rnd in range(8):
            outs = [np.zeros((1, S, H), np.float32) for _ in range(4)]
            res = srv.wait(srv.submit_batch([(i, 1, [ids[i:i + 1].copy(), mask[i:i + 1].copy()], [outs[i]])
                                             for i in range(4)], fixed_worker=rnd % 2))
            assert all(r.ok == 1 for r in res), res[0].error
            devices.add(res[0].device_id)
            assert (np.concatenate(outs, 0) == want).all(), f"round {rnd} on device {res[0].device_id}"
        assert devices == {0, 1}
    for r in reps:
        r.close()



@pytest.mark.parametrize("variable_length", [False, True])
def test_device_visible_requests_through_the_task_path(host, variable_length):
    """SURVEY 8(f) row 2 on the SERVED path (spu_server_cfg.device_visible_requests): requests and responses live in
    pinned host memory, the host gathers nothing -- the input slot carries a table of row addresses -- and the
    codelet runs gather/pad/mask -> encoder -> per-request scatter on the GPU.  Every response is checked against
    the fp32 ORACLE run on that request alone; multi-sample requests, ragged lengths (variable_length) and the
    dynamic-batching collector are covered; the output slot carries no payload (first_value is not the response)."""
    from oracle import bert_oracle as bo
    from starpu_inference_server_b200 import capi
    from starpu_inference_server_b200.bert import BertReplica

    spec = bo.BERT_TINY
    w = bo.make_weights(spec, 0)
    S, H = 32, spec.hidden
    # auto-graph: from its second task of a shape on, a worker replays gather -> encoder -> scatter as one CUDA graph
    # whose table addresses (they move with the slot) are patched per launch
    rep = BertReplica(w, max_batch=8, max_seq=S, n_workspaces=2, flags=capi.SPU_REQUEST_IO | capi.SPU_FORWARD_AUTO_GRAPH)
    rng = np.random.default_rng(33)
    shapes = [(1, 17), (2, 32), (1, 5), (1, 32), (2, 9)] if variable_length else [(1, S), (2, S), (1, S), (3, S)]
    reqs = []
    for b, n in shapes:
        ids = torch.from_numpy(rng.integers(1, spec.vocab, size=(b, n), dtype=np.int64)).pin_memory()
        mask = torch.ones(b, n, dtype=torch.int64)
        if not variable_length:
            for r in range(b):
                mask[r, int(rng.integers(3, n + 1)):] = 0
        mask = mask.pin_memory()
        out = torch.full((b, n, H), float("nan")).pin_memory()
        reqs.append((b, n, ids, mask, out))
    with host.Server(inputs=[("input_ids", "i64", [1, S]), ("attention_mask", "i64", [1, S])],
                     outputs=[("output0", "f32", [1, S, H])], replicas=[rep], model_kind=host.MODEL_BERT,
                     max_batch_size=8, pool_size=2, n_workspaces=2, variable_length=variable_length,
                     device_visible_requests=True, batch_coalesce_timeout_ms=20) as srv:
        def check(tag):
            for i, (b, n, ids, mask, out) in enumerate(reqs):
                ref = bo.bert_forward(w, spec, ids.numpy(), mask.numpy())
                got = out.numpy()
                live = mask.numpy().astype(bool)
                assert np.isfinite(got[live]).all(), (tag, i)
                assert np.abs(got[live] - ref[live]).max() / np.abs(ref).max() < 2e-2, (tag, i)
                out.fill_(float("nan"))

        for rnd in range(5):  # eager, capture, then retargeted replays on the other slot
            res = srv.wait(srv.submit_batch([(i, b, [ids, mask], [out]) for i, (b, n, ids, mask, out) in enumerate(reqs)],
                                            fixed_worker=0))
            assert all(r.ok == 1 for r in res), [r.error for r in res]
            assert res[0].batch_rows == sum(b for b, *_ in reqs) and [r.rows for r in res] == [b for b, *_ in reqs]
            check(f"one task, round {rnd}")
        tickets = [srv.enqueue(100 + i, b, [ids, mask], [out]) for i, (b, n, ids, mask, out) in enumerate(reqs)]
        rows = 0
        for t in tickets:
            (r,) = srv.wait(t)
            assert r.ok == 1, r.error
            rows = max(rows, r.batch_rows)
        assert rows > 1  # the collector merged requests
        check("collector")
        # a request without a response buffer cannot be served on this path
        b, n, ids, mask, out = reqs[0]
        with pytest.raises(host.HostError, match="output buffer"):
            srv.wait(srv.submit_batch([(0, b, [ids, mask], [None])]))
        assert srv.stats()["codelet_failures"] == 0
    rep.close()


def test_device_visible_requests_with_a_token_id_outside_the_table_fail_the_batch(host):
    """on the device_visible_requests path the host never reads the ids, so the gather kernel checks them
    and the output callback fails the batch (ok = 0 for every request of it, as the reference's job fails when
    LibTorch's embedding raises) instead of returning a plausible response computed from a clamped id.  The next
    clean batch on the same worker succeeds: the error word does not stick."""
    from oracle import bert_oracle as bo
    from starpu_inference_server_b200 import capi
    from starpu_inference_server_b200.bert import BertReplica

    spec = bo.BERT_TINY
    w = bo.make_weights(spec, 0)
    S, H = 32, spec.hidden
    rep = BertReplica(w, max_batch=8, max_seq=S, n_workspaces=2, flags=capi.SPU_REQUEST_IO | capi.SPU_FORWARD_AUTO_GRAPH)
    rng = np.random.default_rng(34)
    reqs = []
    for b in (1, 2, 1):
        ids = torch.from_numpy(rng.integers(1, spec.vocab, size=(b, S), dtype=np.int64)).pin_memory()
        mask = torch.ones(b, S, dtype=torch.int64).pin_memory()
        out = torch.full((b, S, H), float("nan")).pin_memory()
        reqs.append((b, ids, mask, out))
    batch = lambda: [(i, b, [ids, mask], [out]) for i, (b, ids, mask, out) in enumerate(reqs)]  # noqa: E731
    with host.Server(inputs=[("input_ids", "i64", [1, S]), ("attention_mask", "i64", [1, S])],
                     outputs=[("output0", "f32", [1, S, H])], replicas=[rep], model_kind=host.MODEL_BERT,
                     max_batch_size=8, pool_size=2, n_workspaces=2, device_visible_requests=True) as srv:
        for rnd in range(4):  # eager, captured, replayed: the verdict follows the task, not the graph
            res = srv.wait(srv.submit_batch(batch(), fixed_worker=0))
            assert all(r.ok == 1 for r in res), [r.error for r in res]
            keep = reqs[1][1][1, 5].item()
            reqs[1][1][1, 5] = spec.vocab  # second sample of request 1 = batch row 2
            res = srv.wait(srv.submit_batch(batch(), fixed_worker=0))
            assert all(r.ok == 0 for r in res)
            assert all(b"outside the embedding table" in r.error and b"row 2" in r.error for r in res), [r.error for r in res]
            reqs[1][1][1, 5] = keep
        res = srv.wait(srv.submit_batch(batch(), fixed_worker=0))
        assert all(r.ok == 1 for r in res)
        ref = bo.bert_forward(w, spec, reqs[1][1].numpy(), reqs[1][2].numpy())
        assert np.abs(reqs[1][3].numpy() - ref).max() / np.abs(ref).max() < 2e-2
    rep.close()


def test_out_of_range_token_id_fails_the_batch_like_the_reference(host, bert_tiny):
    """LibTorch's embedding raises on an id outside the table and the reference's job fails; the kernel clamps (memory
    safety), so the host gather refuses the batch before anything is submitted."""
    bo, spec, w, rep = bert_tiny
    S, H = 32, spec.hidden
    ids, mask = bo.make_inputs(spec, 2, S, seed=3)
    with host.Server(inputs=[("input_ids", "i64", [1, S]), ("attention_mask", "i64", [1, S])],
                     outputs=[("output0", "f32", [1, S, H])], replicas=[rep], model_kind=host.MODEL_BERT,
                     max_batch_size=4, pool_size=2, n_workspaces=2) as srv:
        for bad in (spec.vocab, -1):
            b = ids[1:2].copy()
            b[0, 7] = bad
            with pytest.raises(host.HostError, match="outside the embedding table"):
                srv.submit_batch([(0, 1, [ids[:1].copy(), mask[:1].copy()], [None]), (1, 1, [b, mask[1:2].copy()], [None])])
        out = np.zeros((1, S, H), np.float32)
        (r,) = srv.wait(srv.submit_batch([(0, 1, [ids[:1].copy(), mask[:1].copy()], [out])]))
        assert r.ok == 1 and srv.stats()["free_in"] == 2  # the refused batches gave their slots back
