"""spu_bert_forward_static_async: the forward for callers that cannot obtain device memory.

Once a server is up, code on its serving path may hold a replica and its own input / output buffers but must not
call cudaMalloc any more: inside a StarPU codelet an allocation stalls every other worker's stream on that GPU and
bypasses StarPU's memory accounting, and while any stream is being captured into a CUDA graph it is an error.  A
worker the replica has no workspace for (or a health probe, or a late warm-up) is in that position.  The entry
point tested here runs the encoder on a scratch array that is part of the library image on each GPU
(SPU_STATIC_SCRATCH_BYTES), so the call allocates nothing.

The check is against the existing entry point, spu_bert_forward_async on a replica workspace, through the C-ABI:
same replica, same inputs.  Both enqueue the same kernels with the same tiles, so the outputs must be bit-equal
(asserted), which is inside the 2e-2 relative tolerance the forward tests use (asserted too, so a future change of
tile choice shows up as "still correct, no longer identical" rather than as one opaque failure).  Two calls run
back to back with different shapes and different inputs: the array is shared and keeps the first call's
activations, and the second result must not depend on them.  The file sorts last on purpose (pytest -x).
"""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu
BF16_TOL = 2e-2  # tests/test_gpu_forward.py


def _inputs(cfg, batch, seq, seed):
    rng = np.random.default_rng(seed)
    ids = rng.integers(0, cfg.vocab, size=(batch, seq), dtype=np.int64)
    mask = np.ones((batch, seq), np.int64)
    for b in range(batch):  # ragged right padding, at least one live token per row
        mask[b, int(rng.integers(1, seq + 1)):] = 0
    return torch.from_numpy(ids).cuda(), torch.from_numpy(mask).cuda()


@pytest.fixture(scope="module")
def replica(built_libs):
    from oracle import bert_oracle as bo  # seeded weights only; the oracle's forward is not used here
    from starpu_inference_server_b200 import capi
    from starpu_inference_server_b200.bert import BertReplica

    assert torch.cuda.is_available()
    assert capi.lib().spu_device_is_sm100(0) == 1
    r = BertReplica(bo.make_weights(bo.BERT_BASE, 3), max_batch=4, max_seq=128, n_workspaces=1)
    yield capi, bo.BERT_BASE, r
    r.close()


def _both(r, hidden, ids, mask):
    want = torch.full((*ids.shape, hidden), float("nan"), device="cuda")
    got = torch.full_like(want, float("nan"))
    r.forward_async(ids, mask, want)       # existing entry point, replica workspace 0
    r.forward_static_async(ids, mask, got)  # new entry point, library-held scratch
    torch.cuda.synchronize()
    return got.cpu().numpy(), want.cpu().numpy()


def test_static_scratch_forward_matches_workspace_forward_twice_in_a_row(replica):
    capi, cfg, r = replica
    hidden = cfg.hidden
    # call 1: a small batch; call 2: the largest BERT-base batch the array holds (512 tokens), other inputs
    for batch, seq, seed in ((2, 64, 11), (4, 128, 12)):
        ids, mask = _inputs(cfg, batch, seq, seed)
        got, want = _both(r, hidden, ids, mask)
        assert np.isfinite(got).all()
        assert float(np.abs(got - want).max() / np.abs(want).max()) <= BF16_TOL
        assert np.array_equal(got, want), (batch, seq)
    # and the small shape once more, after the large call has overwritten every byte the small one uses
    ids, mask = _inputs(cfg, 2, 64, 11)
    got, want = _both(r, hidden, ids, mask)
    assert np.array_equal(got, want)


def test_static_scratch_forward_rejects_what_does_not_fit(replica):
    capi, cfg, r = replica
    L = capi.lib()
    ids, mask = _inputs(cfg, 5, 128, 13)  # 640 tokens of BERT-base > 9 MiB of activations
    out = torch.zeros((5, 128, cfg.hidden), device="cuda")
    st = L.spu_bert_forward_static_async(r._h, capi.ptr(ids), capi.ptr(mask), 5, 128, capi.ptr(out),
                                         capi.stream_ptr(None))
    assert st == capi.SPU_ERR_CAPACITY
    msg = L.spu_last_error().decode()
    assert "SPU_STATIC_SCRATCH_BYTES" in msg and str(capi.SPU_STATIC_SCRATCH_BYTES) in msg
    torch.cuda.synchronize()
    assert float(out.abs().max()) == 0.0  # nothing was enqueued
    st = L.spu_bert_forward_static_async(r._h, None, capi.ptr(mask), 5, 128, capi.ptr(out), capi.stream_ptr(None))
    assert st == capi.SPU_ERR_INVALID
