#!/usr/bin/env python
"""GPU-box tool: fused attention, tcgen05/TMEM kernel (impl 1) against the mma.sync kernel (impl 2) at the
shapes of BASELINE.json's configs.  Each launch works on its own qkv/ctx pair out of a ring larger than
L2 unless --warm (the in-situ case: qkv was just written by the QKV GEMM and is L2-resident).
Reports us per launch, achieved GB/s on the ALGORITHMIC bytes (read q,k,v + write ctx) against the
measured copy bandwidth, and TFLOP/s on 4*S*S*64 per (batch, head)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from starpu_inference_server_b200 import capi  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=50)
    ap.add_argument("--warm", action="store_true")
    args = ap.parse_args()
    L = capi.lib()
    peak = 6553.9
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = float(json.load(open(p))["hbm_gbs"])
    shapes = [("bert b32 s128", 32, 128, 12), ("bert b256 s128", 256, 128, 12), ("bert b64 s512", 64, 512, 12),
              ("bert b64 s384", 64, 384, 12), ("vit-l b64 s197", 64, 197, 16), ("bert b1 s128", 1, 128, 12)]
    s = torch.cuda.Stream()
    sp = s.cuda_stream
    for name, B, S, heads in shapes:
        H = heads * 64
        rows = B * S
        nbytes = rows * H * 2 * 4
        ring = 1 if args.warm else max(2, int(300e6 // nbytes) + 1)
        qkv = [torch.randn(rows, 3 * H, device="cuda").to(torch.bfloat16) for _ in range(ring)]
        ctx = [torch.empty(rows, H, dtype=torch.bfloat16, device="cuda") for _ in range(ring)]
        mask = torch.zeros(B, S, device="cuda")
        flops = 4.0 * S * S * 64 * heads * B
        for impl, iname in ((1, "tcgen05"), (2, "mma.sync")):
            def run(i):
                st = L.spu_attention_cfg_async(qkv[i % ring].data_ptr(), mask.data_ptr(), ctx[i % ring].data_ptr(), B, S,
                                               heads, impl, 0, 1.0, sp)
                assert st == 0, L.spu_last_error()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with torch.cuda.stream(s):
                for i in range(3):
                    run(i)
                e0.record(s)
                for i in range(args.reps):
                    run(i)
                e1.record(s)
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / args.reps
            gbs = nbytes / (ms / 1e3) / 1e9
            print(json.dumps({"shape": name, "impl": iname, "us": round(ms * 1e3, 2), "GB/s": round(gbs, 1),
                              "frac_of_hbm_peak": round(gbs / peak, 3), "TFLOP/s": round(flops / (ms / 1e3) / 1e12, 1),
                              "l2": "warm" if args.warm else f"ring of {ring} > L2"}), flush=True)
        del qkv, ctx


if __name__ == "__main__":
    main()
