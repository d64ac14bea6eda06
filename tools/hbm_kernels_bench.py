#!/usr/bin/env python
"""GPU-box tool: achieved HBM GB/s of the memory-bound kernels of the path (gather/pad/mask,
embedding+LN, residual LayerNorm, fused attention, scatter) at BERT-base shapes, against the
measured copy bandwidth in MEASURED_PEAKS.json.  Each kernel runs `reps` times back to back on
buffers larger than L2 (126 MB) where the shape allows, CUDA events on the launch stream;
bytes are the ALGORITHMIC bytes of DESIGN.md section 4."""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from starpu_inference_server_b200 import capi  # noqa: E402


def timed(fn, reps):
    s = torch.cuda.Stream()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(s):
        for _ in range(3):
            fn(s)
        e0.record(s)
        for _ in range(reps):
            fn(s)
        e1.record(s)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def table(ptrs):
    return torch.tensor(ptrs, dtype=torch.int64, device="cuda")


def main():
    L = capi.lib()
    peak = 6453.4
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = float(json.load(open(p))["hbm_gbs"])
    B, S, H, heads = 256, 128, 768, 12
    rows = B * S
    out = []

    def rec(name, nbytes, ms, note=""):
        gbs = nbytes / (ms / 1e3) / 1e9
        out.append({"kernel": name, "bytes": nbytes, "us": ms * 1e3, "GB/s": gbs, "frac_of_peak": gbs / peak, "note": note})

    # ---- stage 1: ragged gather + pad + mask, 8192 requests of <= 512 tokens --------------
    n_req, seq = 8192, 512
    rng = np.random.default_rng(0)
    lens = rng.integers(16, seq + 1, size=n_req).astype(np.int32)
    src = torch.randint(0, 30522, (int(lens.sum()),), dtype=torch.int64, device="cuda")
    offs = np.concatenate([[0], np.cumsum(lens)[:-1]])
    ptrs = table([src.data_ptr() + 8 * int(o) for o in offs])
    lens_d = torch.from_numpy(lens).cuda()
    ids_o = torch.empty(n_req, seq, dtype=torch.int64, device="cuda")
    mask_o = torch.empty_like(ids_o)
    ms = timed(lambda s: capi.check(L.spu_gather_pad_mask_async(ptrs.data_ptr(), None, lens_d.data_ptr(), n_req, seq, 0,
                                                                ids_o.data_ptr(), mask_o.data_ptr(),
                                                                C.c_void_p(s.cuda_stream))), 20)
    rec("gather_pad_mask_kernel", int(lens.sum()) * 8 + 2 * n_req * seq * 8, ms, "8192 ragged requests -> [8192,512] ids+mask")

    # ---- embedding + LN + mask bias ---------------------------------------------------------
    word = torch.randn(30522, H, device="cuda")
    pos = torch.randn(512, H, device="cuda")
    typ = torch.randn(H, device="cuda")
    g, b = torch.ones(H, device="cuda"), torch.zeros(H, device="cuda")
    ids = torch.randint(0, 30522, (B, S), dtype=torch.int64, device="cuda")
    mask = torch.ones_like(ids)
    x32 = torch.empty(rows, H, device="cuda")
    x16 = torch.empty(rows, H, dtype=torch.bfloat16, device="cuda")
    mb = torch.empty(rows, device="cuda")
    ms = timed(lambda s: capi.check(L.spu_bert_embed_async(ids.data_ptr(), mask.data_ptr(), word.data_ptr(), pos.data_ptr(),
                                                           typ.data_ptr(), g.data_ptr(), b.data_ptr(), 1e-12, 30522,
                                                           x32.data_ptr(), x16.data_ptr(), mb.data_ptr(), B, S, H,
                                                           C.c_void_p(s.cuda_stream))), 20)
    rec("bert_embed_ln_kernel", rows * H * (4 + 4 + 2) + rows * 16, ms, "word rows read fp32 (pos table L2-resident), x fp32+bf16 written")

    # ---- residual LayerNorm -----------------------------------------------------------------
    xin = torch.randn(rows, H, device="cuda")
    ms = timed(lambda s: capi.check(L.spu_layernorm_async(xin.data_ptr(), g.data_ptr(), b.data_ptr(), 1e-12, x32.data_ptr(),
                                                          x16.data_ptr(), rows, H, C.c_void_p(s.cuda_stream))), 20)
    rec("layernorm_kernel", rows * H * (4 + 4 + 2), ms, "[32768,768] fp32 in, fp32 + bf16 out")

    # ---- fused attention --------------------------------------------------------------------
    qkv = (torch.randn(rows, 3 * H, device="cuda") * 0.5).to(torch.bfloat16)
    ctx = torch.empty(rows, H, dtype=torch.bfloat16, device="cuda")
    bias = torch.zeros(B, S, device="cuda")
    ms = timed(lambda s: capi.check(L.spu_attention_async(qkv.data_ptr(), bias.data_ptr(), ctx.data_ptr(), B, S, heads,
                                                          C.c_void_p(s.cuda_stream))), 20)
    rec("attention_kernel", rows * H * 2 * 4 + rows * 4, ms, "qkv read once + ctx written, B=256 S=128")

    # ---- stage 4: scatter [B,S,H] fp32 rows to per-request buffers --------------------------
    n = 512
    big = torch.randn(n, S, H, device="cuda")
    dsts = [torch.empty(S * H, device="cuda") for _ in range(n)]
    dptr = table([d.data_ptr() for d in dsts])
    soff = torch.arange(n, dtype=torch.int64, device="cuda") * (S * H * 4)
    nby = torch.full((n,), S * H * 4, dtype=torch.int64, device="cuda")
    ms = timed(lambda s: capi.check(L.spu_scatter_segments_async(big.data_ptr(), soff.data_ptr(), nby.data_ptr(),
                                                                 dptr.data_ptr(), n, C.c_void_p(s.cuda_stream))), 20)
    rec("segcopy_kernel (scatter)", 2 * n * S * H * 4, ms, "512 requests x 393216 B")
    ok = all(torch.equal(d.view(S, H), big[i]) for i, d in enumerate(dsts))
    assert ok
    for r in out:
        print(json.dumps(r))


if __name__ == "__main__":
    main()
