#!/usr/bin/env python
"""Tile sweep of the tcgen05 GEMM on the BERT-base / ViT-L shapes (GPU box only).

Each (shape, epilogue, cta_group, BLOCK_N) runs `reps` back-to-back launches captured in
one CUDA graph (so host encode/ctypes time is excluded and programmatic dependent launch
is in effect, as in the replica's captured forward) and is timed with CUDA events.
Prints one JSON line per configuration; correctness is tests/test_gpu_stages.py's job.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from starpu_inference_server_b200 import capi  # noqa: E402

SHAPES = {  # name: (N, K, epilogue)
    "qkv": (2304, 768, 0), "attn_out": (768, 768, 2), "ffn1": (3072, 768, 1), "ffn2": (768, 3072, 2),
    # epilogue-cost probes: the N=768 shapes with the plain bf16 epilogue / with a separate residual buffer
    # residual + LayerNorm fused into the epilogue (cluster of 4 CTA pairs x 192 columns)
    "attn_out_ln": (768, 768, 5), "ffn2_ln": (768, 3072, 5),
    "attn_out_bf16": (768, 768, 0), "ffn2_bf16": (768, 3072, 0), "attn_out_sep": (768, 768, 12), "ffn2_sep": (768, 3072, 12),
}


def run(M, N, K, epi, cg, bn, reps, n_buf, n_streams=1):
    if n_streams > 1:
        return run_streams(M, N, K, epi, cg, bn, reps, n_streams)
    L = capi.lib()
    separate = epi == 12  # BIAS_RESID reading the residual from a different buffer (no in-place reduce-add)
    epi = 2 if separate else epi
    a = [(torch.randn(M, K, device="cuda") * 0.5).to(torch.bfloat16) for _ in range(n_buf)]
    w = (torch.randn(N, K, device="cuda") * 0.05).to(torch.bfloat16)
    bias = torch.randn(N, device="cuda")
    out = [torch.zeros(M, N, device="cuda", dtype=torch.float32 if epi == 2 else torch.bfloat16) for _ in range(n_buf)]
    res = [torch.zeros(M, N, device="cuda") for _ in range(n_buf)] if separate else out
    s = torch.cuda.Stream()

    ln = epi == 5
    if ln:
        out = [torch.zeros(M, N, device="cuda") for _ in range(n_buf)]
        out16 = [torch.zeros(M, N, device="cuda", dtype=torch.bfloat16) for _ in range(n_buf)]
        gamma, beta = torch.ones(N, device="cuda"), torch.zeros(N, device="cuda")

    def enqueue():
        for i in range(reps):
            o = out[i % n_buf]
            if ln:
                capi.check(L.spu_gemm_ln_bf16_async(capi.ptr(a[i % n_buf]), capi.ptr(w), capi.ptr(bias), capi.ptr(o),
                                                    capi.ptr(gamma), capi.ptr(beta), 1e-12, capi.ptr(o),
                                                    capi.ptr(out16[i % n_buf]), M, N, K, cg, bn, capi.stream_ptr(s)))
                continue
            capi.check(L.spu_gemm_bf16_cfg_async(capi.ptr(a[i % n_buf]), capi.ptr(w), capi.ptr(bias),
                                                 capi.ptr(res[i % n_buf]) if epi == 2 else None, capi.ptr(o), M, N, K, epi,
                                                 cg, bn, capi.stream_ptr(s)))
    with torch.cuda.stream(s):
        enqueue()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        enqueue()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for _ in range(5):
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / reps * 1e3)
    return best


def run_streams(M, N, K, epi, cg, bn, reps, n_streams):
    """`n_streams` CUDA workers, each replaying its own graph of `reps` launches on private buffers, all at once:
    microseconds per launch of the whole device (elapsed / (reps * n_streams)) -- what co-resident CTAs of two
    in-flight batches buy.  epi 2 = in-place residual accumulate as in the forward."""
    L = capi.lib()
    main_s = torch.cuda.Stream()
    graphs, streams, keep = [], [], []
    for _ in range(n_streams):
        a = (torch.randn(M, K, device="cuda") * 0.5).to(torch.bfloat16)
        w = (torch.randn(N, K, device="cuda") * 0.05).to(torch.bfloat16)
        bias = torch.randn(N, device="cuda")
        out = torch.zeros(M, N, device="cuda", dtype=torch.float32 if epi == 2 else torch.bfloat16)
        keep += [a, w, bias, out]
        s = torch.cuda.Stream()

        def enqueue(a=a, w=w, bias=bias, out=out, s=s):
            for _ in range(reps):
                capi.check(L.spu_gemm_bf16_cfg_async(capi.ptr(a), capi.ptr(w), capi.ptr(bias),
                                                     capi.ptr(out) if epi == 2 else None, capi.ptr(out), M, N, K, epi,
                                                     cg, bn, capi.stream_ptr(s)))
        with torch.cuda.stream(s):
            enqueue()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            enqueue()
        graphs.append(g)
        streams.append(s)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for _ in range(5):
        with torch.cuda.stream(main_s):
            e0.record(main_s)
            for g, s in zip(graphs, streams):
                s.wait_event(e0)
                with torch.cuda.stream(s):
                    g.replay()
                ev = torch.cuda.Event()
                ev.record(s)
                main_s.wait_event(ev)
            e1.record(main_s)
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / (reps * n_streams) * 1e3)
    return best


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--M", default="4096", help="comma-separated row counts")
    ap.add_argument("--reps", type=int, default=40)
    ap.add_argument("--bufs", type=int, default=1, help="distinct A/out buffers rotated (1 = L2-warm, as in the forward)")
    ap.add_argument("--shapes", default="qkv,attn_out,ffn1,ffn2")
    ap.add_argument("--all", action="store_true", help="sweep every tile, not only the chosen one")
    ap.add_argument("--cfgs", default="", help="explicit tiles 'cg:bn,...' (cg 3 = CTA pair of half-footprint CTAs)")
    ap.add_argument("--streams", type=int, default=1, help="concurrent CUDA workers, each with its own graph + buffers")
    args = ap.parse_args()
    for name, M in ((n, int(m)) for n in args.shapes.split(",") for m in args.M.split(",")):
        N, K, epi = SHAPES[name]
        cfgs = [(0, 0)]
        if args.all and epi != 5:
            cfgs += [(cg, bn) for cg in (2, 1) for bn in (256, 192, 128, 64) if N % bn == 0]
        if args.cfgs:
            cfgs = [tuple(int(v) for v in c.split(":")) for c in args.cfgs.split(",")]
        for cg, bn in cfgs:
            try:
                us = run(M, N, K, epi, cg, bn, args.reps, args.bufs, args.streams)
            except Exception as ex:  # keep sweeping
                print(json.dumps({"shape": name, "cg": cg, "bn": bn, "error": str(ex)}), flush=True)
                continue
            tf = 2.0 * M * N * K / us / 1e6
            print(json.dumps({"shape": name, "M": M, "N": N, "K": K, "epi": epi, "cg": cg, "bn": bn,
                              "streams": args.streams, "us": round(us, 2), "tflops": round(tf, 1)}), flush=True)


if __name__ == "__main__":
    main()
