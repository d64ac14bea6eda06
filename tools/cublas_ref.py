#!/usr/bin/env python
"""GPU-box tool: the library baseline for the encoder GEMMs.  Times cuBLAS (torch.addmm, bf16, bias only -- no
GELU, no residual, no fused store format) on the four GEMM shapes of one BERT-base layer at M = batch*seq rows,
as a 12-layer chain of 48 launches captured in one CUDA graph, exactly like bench.py times this repo's
gemm_tcgen05_kernel chain (`roofline.kernel_ms_per_step`).  Prints one JSON line."""
import json
import sys

import torch

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
S = int(sys.argv[2]) if len(sys.argv) > 2 else 128
M, H, F, LAYERS = B * S, 768, 3072, 12
shapes = {"qkv": (3 * H, H), "attn_out": (H, H), "ffn1": (F, H), "ffn2": (H, F)}
dev = "cuda"
w = {k: [(torch.randn(n, kk, device=dev) * 0.02).to(torch.bfloat16) for _ in range(LAYERS)] for k, (n, kk) in shapes.items()}
bias = {k: torch.randn(n, device=dev).to(torch.bfloat16) for k, (n, kk) in shapes.items()}
x = {k: (torch.randn(M, kk, device=dev) * 0.5).to(torch.bfloat16) for k, (n, kk) in shapes.items()}
out = {k: torch.empty(M, n, device=dev, dtype=torch.bfloat16) for k, (n, kk) in shapes.items()}


def chain(names, forwards=5):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(2):
            for name in names:
                torch.addmm(bias[name], x[name], w[name][0].t(), out=out[name])
        s.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(forwards):
                for layer in range(LAYERS):
                    for name in names:
                        torch.addmm(bias[name], x[name], w[name][layer].t(), out=out[name])
        best = 1e9
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            g.replay()
            e0.record(s)
            g.replay()
            e1.record(s)
            s.synchronize()
            best = min(best, e0.elapsed_time(e1) / forwards)
    return best


res = {"what": f"cuBLAS bf16 addmm chain, M={M}", "ms_per_forward": {}}
for name in shapes:
    res["ms_per_forward"][name] = chain([name])
res["ms_per_forward"]["all_48"] = chain(list(shapes))
flop = LAYERS * 24 * M * H * H
res["tflops_all_48"] = flop / (res["ms_per_forward"]["all_48"] / 1e3) / 1e12
print(json.dumps(res))
