#!/usr/bin/env python
"""GPU-box tool: SURVEY 8(f) rows 1+2 measured at the codelet level.  Requests ([S] ids + mask each) and responses
([S, H] fp32 each) live in PINNED HOST memory; spu_bert_forward_requests_async reads the requests over PCIe in its
gather/pad/mask kernel, runs the encoder, and its scatter kernel writes every response straight into that request's
own host buffer -- no host gather, no H2D/D2H staging copy, no slot slicing.  Reports inferences/s over `steps`
batches issued round-robin on `workers` streams (one workspace each), host wall clock around the whole loop."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from starpu_inference_server_b200 import capi, synthetic  # noqa: E402
from starpu_inference_server_b200.bert import BertConfig, BertReplica  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--seq", type=int, default=128)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--workers", type=int, default=4)
    ap.add_argument("--slots", type=int, default=12)
    ap.add_argument("--min-len", type=int, default=None)
    ap.add_argument("--live-rows", action="store_true", help="responses carry only the live rows of each request")
    args = ap.parse_args()
    B, S = args.batch, args.seq
    spec = BertConfig()
    rep = BertReplica(synthetic.make_weights(spec, 0), max_batch=B, max_seq=S, n_workspaces=args.workers,
                      flags=capi.SPU_REQUEST_IO)
    H = spec.hidden
    rng = np.random.default_rng(0)
    slots = []
    for k in range(args.slots):
        ids, mask = synthetic.make_requests(spec, B, S, seed=100 + k, min_len=args.min_len)
        lens = mask.sum(1).astype(np.int32)
        rid = [torch.from_numpy(np.ascontiguousarray(ids[i, :lens[i]])).pin_memory() for i in range(B)]
        rout = [torch.empty((int(lens[i]) if args.live_rows else S), H).pin_memory() for i in range(B)]
        slots.append({
            "ids": rid, "out": rout,
            "t_ids": torch.tensor([t.data_ptr() for t in rid], dtype=torch.int64, device="cuda"),
            "t_out": torch.tensor([t.data_ptr() for t in rout], dtype=torch.int64, device="cuda"),
            "len": torch.from_numpy(lens).cuda(), "ev": torch.cuda.Event(), "busy": False,
            "bytes_out": sum(t.numel() * 4 for t in rout), "bytes_in": int(lens.sum()) * 8})
    streams = [torch.cuda.Stream() for _ in range(args.workers)]
    torch.cuda.synchronize()

    def run(n):
        checksum = 0.0
        for i in range(n):
            sl = slots[i % args.slots]
            if sl["busy"]:
                sl["ev"].synchronize()          # response buffers of this slot are complete in host memory
                checksum += float(sl["out"][0][0, 0])
            w = i % args.workers
            rep.forward_requests_async(sl["t_ids"], None, sl["len"], B, S, sl["t_out"], full_rows=not args.live_rows,
                                       ws=w, stream=streams[w])
            sl["ev"].record(streams[w])
            sl["busy"] = True
        for sl in slots:
            if sl["busy"]:
                sl["ev"].synchronize()
                sl["busy"] = False
        return checksum

    run(3 * args.slots)
    t0 = time.perf_counter()
    cs = run(args.steps)
    dt = time.perf_counter() - t0
    print(json.dumps({"workload": f"BERT-base b{B} s{S} requests/responses in pinned host memory, zero staging copies"
                                  + (f", lengths U[{args.min_len},{S}]" if args.min_len else ""),
                      "inferences_per_s": B * args.steps / dt, "ms_per_batch": dt / args.steps * 1e3,
                      "workers": args.workers, "slots": args.slots, "live_rows_only": args.live_rows,
                      "host_read_bytes_per_batch": slots[0]["bytes_in"],
                      "host_write_bytes_per_batch": slots[0]["bytes_out"], "checksum": cs}))
    rep.close()


if __name__ == "__main__":
    main()
