#!/usr/bin/env python
"""GPU-box tool: per-kernel-class chain times (CUDA graph, events) of the BERT forward."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from starpu_inference_server_b200 import capi, synthetic  # noqa: E402
from starpu_inference_server_b200.bert import BertConfig, BertReplica  # noqa: E402


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
    S = int(sys.argv[2]) if len(sys.argv) > 2 else 128
    spec = BertConfig()
    fp8 = len(sys.argv) > 3 and sys.argv[3].startswith("fp8")
    flags = 0
    if fp8:
        n8 = int(sys.argv[3][3:] or 0)
        flags = capi.SPU_PREC_FP8_E4M3 | (n8 << 8)
    rep = BertReplica(synthetic.make_weights(spec, 0), max_batch=B, max_seq=S, flags=flags)
    ids, mask = synthetic.make_requests(spec, B, S, seed=1)
    ids, mask = torch.from_numpy(ids).cuda(), torch.from_numpy(mask).cuda()
    if fp8:
        rep.calibrate_fp8(ids[:32].contiguous(), mask[:32].contiguous())
    out = torch.empty(B, S, spec.hidden, device="cuda")
    res = {"B": B, "S": S, "prec": sys.argv[3] if fp8 else "bf16"}
    st = torch.cuda.Stream()
    for name in capi.STAGE_NAMES:
        res[name] = min(rep.stage_chain_ms(ids, mask, out, [name], stream=st) for _ in range(3))
    res["gemms"] = min(rep.stage_chain_ms(ids, mask, out, [n for n in capi.STAGE_NAMES if n.startswith("gemm")],
                                          stream=st)
                       for _ in range(3))
    res["all_graph"] = min(rep.stage_chain_ms(ids, mask, out, None, stream=st) for _ in range(3))
    s = torch.cuda.Stream()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(s):
        for _ in range(5):
            rep.forward_async(ids, mask, out, stream=s)
        e0.record(s)
        for _ in range(50):
            rep.forward_async(ids, mask, out, stream=s)
        e1.record(s)
    torch.cuda.synchronize()
    res["eager"] = e0.elapsed_time(e1) / 50
    flops = 12 * 24 * S * 768 * 768 * B
    res["gemm_tflops_chain"] = flops / (res["gemms"] / 1e3) / 1e12
    res["sum_classes"] = sum(res[n] for n in capi.STAGE_NAMES)
    print(json.dumps(res))
    rep.close()


if __name__ == "__main__":
    main()
