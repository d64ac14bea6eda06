#!/usr/bin/env python
"""GPU-box tool: isolate the MMA rate from the operand-feed rate of the tcgen05 GEMM.

For each (shape, cta_group, BLOCK_N) it times `reps` back-to-back launches in a CUDA graph in
three modes (spu_debug_gemm_mode): normal, MMA only (no TMA operand loads) and feed only (no
MMAs).  Results of modes 1/2 are garbage by construction; only the times mean something."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from starpu_inference_server_b200 import capi  # noqa: E402
from tools.gemm_bench import SHAPES, run  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--M", type=int, default=4096)
    ap.add_argument("--reps", type=int, default=40)
    ap.add_argument("--shapes", default="qkv,ffn2")
    ap.add_argument("--cfgs", default="1:256,2:256,1:192,1:128,2:128")
    ap.add_argument("--stages", default="0", help="comma list of ring-stage limits (0 = all that fit)")
    args = ap.parse_args()
    L = capi.debug_hooks()  # SPU_DEBUG_HOOKS build only
    for name in args.shapes.split(","):
        N, K, epi = SHAPES[name]
        for cfg in args.cfgs.split(","):
            cg, bn = (int(x) for x in cfg.split(":"))
            if N % bn:
                continue
            for st in (int(x) for x in args.stages.split(",")):
                row = {"shape": name, "M": args.M, "N": N, "K": K, "cg": cg, "bn": bn, "stages": st}
                for mode, label in ((0, "normal"), (1, "mma_only"), (2, "feed_only"), (3, "mma2x_only")):
                    L.spu_debug_gemm_mode(mode | (st << 8))
                    row[label + "_us"] = round(run(args.M, N, K, epi, cg, bn, args.reps, 1), 2)
                L.spu_debug_gemm_mode(0)
                tiles = (N // bn) * ((args.M + 128 * cg - 1) // (128 * cg))
                rounds = -(-tiles // (148 // cg))
                row["rounds"] = rounds
                row["kblocks_per_cta"] = rounds * K // 64
                print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
