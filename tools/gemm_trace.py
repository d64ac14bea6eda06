#!/usr/bin/env python
"""GPU-box tool: pipeline time stamps of one CTA of the tcgen05 GEMM (debug trace hook)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from starpu_inference_server_b200 import capi  # noqa: E402

SHAPES = {"qkv": (2304, 768, 0), "attn_out": (768, 768, 2), "ffn1": (3072, 768, 1), "ffn2": (768, 3072, 2),
          "attn_out_ln": (768, 768, 5), "ffn2_ln": (768, 3072, 5)}


TILE = int(sys.argv[1]) if len(sys.argv) > 1 else 0  # local tile whose epilogue is stamped


def main():
    L = capi.debug_hooks()  # SPU_DEBUG_HOOKS build only
    M = int(os.environ.get("TRACE_M", 4096))
    for name, (N, K, epi) in SHAPES.items():
        for cg, bn in ((2, 256 if N != 768 else 192),):
            a = (torch.randn(M, K, device="cuda") * 0.5).to(torch.bfloat16)
            w = (torch.randn(N, K, device="cuda") * 0.05).to(torch.bfloat16)
            bias = torch.randn(N, device="cuda")
            out = torch.zeros(M, N, device="cuda", dtype=torch.float32 if epi == 2 else torch.bfloat16)
            x32 = torch.zeros(M, N, device="cuda")
            tr = torch.zeros(128, dtype=torch.int64, device="cuda")
            s = torch.cuda.Stream()
            with torch.cuda.stream(s):
                for it in range(6):  # last launch is the traced one, after 5 back-to-back warm launches
                    L.spu_debug_gemm_trace(tr.data_ptr() if it == 5 else None, (2 * (cg == 2) + 4) | (TILE << 16))
                    if epi == 5:
                        capi.check(L.spu_gemm_ln_bf16_async(capi.ptr(a), capi.ptr(w), capi.ptr(bias), capi.ptr(x32),
                                                            capi.ptr(bias), capi.ptr(bias), 1e-12, capi.ptr(x32),
                                                            capi.ptr(out), M, N, K, cg, bn, capi.stream_ptr(s)))
                        continue
                    capi.check(L.spu_gemm_bf16_cfg_async(capi.ptr(a), capi.ptr(w), capi.ptr(bias),
                                                         capi.ptr(out) if epi == 2 else None, capi.ptr(out), M, N, K,
                                                         epi, cg, bn, capi.stream_ptr(s)))
            torch.cuda.synchronize()
            L.spu_debug_gemm_trace(None, 0)
            t = tr.cpu().tolist()
            t0 = t[0]
            nkb = K // 64
            kb = [t[2 + i] - t0 for i in range(min(nkb, 60))]
            gaps = [kb[i + 1] - kb[i] for i in range(len(kb) - 1)]
            print(json.dumps({"shape": name, "cg": cg, "bn": bn, "wait_done": t[1] - t0, "first_full": kb[0],
                              "kb_gap_ns_mean": sum(gaps) / max(1, len(gaps)), "kb_gaps_head": gaps[:14],
                              "last_full": kb[-1], "acc_ready": t[62] - t0, "epi_done": t[63] - t0,
                              "exit": t[64] - t0,
                              "ln(ld_landed,pass1,sent,stats,stored0,stored1,stored2)":
                                  [t[70 + i] - t0 for i in range(7)] if epi == 5 else None,
                              "chunks(top,ld_issued,slot_free,ld_landed,staged,fenced,stored,end)":
                                  [[t[66 + 8 * c + i] - t0 for i in range(8)] for c in range(4) if t[66 + 8 * c]]}))


if __name__ == "__main__":
    main()
