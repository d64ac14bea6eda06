"""GPU sweep (checker-side script, not collected by pytest): which of the 48 BERT-base GEMMs can run
on per-tensor-scaled E4M3 operands while the logits stay inside BASELINE's FP8 gate (max relative
error <= 6e-2 of the fp32 oracle, >= 99.5 % top-1)?

    python tests/sweeps/fp8_sweep.py [--out gpurun_out/fp8_sweep.json]

1. error of each (layer, GEMM kind) ALONE in FP8 (variances add to first order);
2. knapsack by noise variance per FLOP moved to FP8, under several variance budgets;
3. every candidate mask is re-measured on fresh seeds at b32 and at config D's b256.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

from oracle import bert_oracle as bo  # noqa: E402
from starpu_inference_server_b200 import capi  # noqa: E402
from starpu_inference_server_b200.bert import BertReplica  # noqa: E402

KINDS = ["qkv", "out", "ffn1", "ffn2"]
FLOPS = {"qkv": 3.0, "out": 1.0, "ffn1": 4.0, "ffn2": 4.0}  # units of 2*M*H*H


def metrics(got, ref):
    d = (got - ref).abs()
    scale = float(ref.abs().max())
    top2 = ref.topk(2, dim=-1).values
    gap = top2[..., 0] - top2[..., 1]
    agree = got.argmax(-1) == ref.argmax(-1)
    decided = gap > 1e-2
    return {"max_rel": float(d.max()) / scale, "rms": float((d * d).mean().sqrt()), "scale": scale,
            "top1_strict": float(agree.float().mean()),
            "top1_margin": float(agree[decided].float().mean()), "n_decided": int(decided.sum()),
            "n_tokens": int(agree.numel())}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "fp8_sweep.json"))
    ap.add_argument("--big", type=int, default=256)
    args = ap.parse_args()
    spec = bo.BERT_BASE
    w = bo.make_weights(spec, 0)
    rep = BertReplica(w, max_batch=args.big, max_seq=128, n_workspaces=1, flags=capi.SPU_PREC_FP8_E4M3)
    ci, cm = bo.make_inputs(spec, 32, 128, seed=999, min_len=16)
    rep.calibrate_fp8(torch.from_numpy(ci).cuda(), torch.from_numpy(cm).cuda())

    def batch(B, seed, min_len=None):
        ids, mask = bo.make_inputs(spec, B, 128, seed=seed, min_len=min_len)
        t0 = time.time()
        ref = torch.from_numpy(bo.bert_forward(w, spec, ids, mask)).cuda()
        print(f"oracle b{B} seed {seed}: {time.time() - t0:.1f} s", flush=True)
        return torch.from_numpy(ids).cuda(), torch.from_numpy(mask).cuda(), ref

    small = [batch(32, 5, 16), batch(32, 6, None)]
    out32 = torch.empty(32, 128, 768, device="cuda")

    def run(mask, data, out):
        rep.set_fp8_mask(mask)
        res = []
        for ids, m, ref in data:
            rep.forward_async(ids, m, out[: ids.shape[0]])
            torch.cuda.synchronize()
            res.append(metrics(out[: ids.shape[0]], ref))
        return {"max_rel": max(r["max_rel"] for r in res), "rms": float(np.mean([r["rms"] for r in res])),
                "top1_strict": min(r["top1_strict"] for r in res), "top1_margin": min(r["top1_margin"] for r in res),
                "scale": res[0]["scale"]}

    L = spec.layers
    rec = {"single": {}, "named": {}, "candidates": []}
    zero = np.zeros(L, dtype=np.uint8)
    rec["named"]["none(bf16)"] = run(zero, small, out32)
    rec["named"]["all"] = run(np.full(L, 15, dtype=np.uint8), small, out32)
    print("bf16:", rec["named"]["none(bf16)"], "\nall fp8:", rec["named"]["all"], flush=True)
    for k, name in enumerate(KINDS):
        rec["named"]["only_" + name] = run(np.full(L, 1 << k, dtype=np.uint8), small, out32)
        rec["named"]["all_but_" + name] = run(np.full(L, 15 & ~(1 << k), dtype=np.uint8), small, out32)
        print(name, "only:", rec["named"]["only_" + name]["max_rel"], rec["named"]["only_" + name]["rms"],
              " all but:", rec["named"]["all_but_" + name]["max_rel"], rec["named"]["all_but_" + name]["rms"], flush=True)
    for n in range(1, L):
        m = zero.copy()
        m[:n] = 15
        rec["named"][f"first_{n}"] = run(m, small, out32)
        m = zero.copy()
        m[L - n:] = 15
        rec["named"][f"last_{n}"] = run(m, small, out32)
    base_var = rec["named"]["none(bf16)"]["rms"] ** 2
    var = np.zeros((L, 4))
    for l in range(L):
        for k in range(4):
            m = zero.copy()
            m[l] = 1 << k
            r = run(m, small, out32)
            rec["single"][f"{l}.{KINDS[k]}"] = r
            var[l, k] = max(r["rms"] ** 2 - base_var, 1e-12)
    print("single-GEMM noise variance (rows = layers; qkv out ffn1 ffn2):\n", np.array2string(var, precision=6), flush=True)
    total_var = float(var.sum())
    print("sum of singles", total_var + base_var, "all-fp8 measured", rec["named"]["all"]["rms"] ** 2, flush=True)
    # knapsack (greedy by variance per FLOP) under variance budgets
    items = sorted(((var[l, k] / FLOPS[KINDS[k]], l, k) for l in range(L) for k in range(4)))
    cands = {}
    for frac in (0.3, 0.4, 0.5, 0.6, 0.7, 0.8):
        budget, acc, m = frac * total_var, 0.0, zero.copy()
        for _, l, k in items:
            if acc + var[l, k] <= budget:
                acc += var[l, k]
                m[l] |= 1 << k
        cands[f"knap_{frac}"] = m
    for n in (6, 7, 8, 9):
        m = zero.copy()
        m[:n] = 15
        cands[f"first_{n}"] = m
    for keep in (3, 7, 11, 13, 14):  # kinds kept everywhere
        cands[f"kinds_{keep}"] = np.full(L, keep, dtype=np.uint8)
    big = [batch(args.big, 7, None)]
    fresh = [batch(32, 11, 16), batch(32, 12, None)]
    outB = torch.empty(args.big, 128, 768, device="cuda")
    for name, m in cands.items():
        flops = sum(FLOPS[KINDS[k]] for l in range(L) for k in range(4) if m[l] >> k & 1) / (12.0 * L)
        r32, rB = run(m, fresh, out32), run(m, big, outB)
        rec["candidates"].append({"name": name, "mask": [int(x) for x in m], "fp8_flop_fraction": flops,
                                  "b32": r32, f"b{args.big}": rB})
        print(f"{name:10s} mask {[int(x) for x in m]} fp8 flops {flops:.2f}  b32 max_rel {r32['max_rel']:.4f} "
              f"top1 {r32['top1_margin']:.4f}/{r32['top1_strict']:.4f}  b{args.big} max_rel {rB['max_rel']:.4f} "
              f"top1 {rB['top1_margin']:.4f}/{rB['top1_strict']:.4f}", flush=True)
    rec["var"] = var.tolist()
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(rec, open(args.out, "w"), indent=1)
    rep.close()


if __name__ == "__main__":
    main()
