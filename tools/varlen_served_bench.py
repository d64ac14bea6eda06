#!/usr/bin/env python
"""GPU-box tool: BASELINE configs[2] (BERT-base, request lengths U[16,512], batch cap 64) through the SERVED path
(spu_server_enqueue: dynamic batching by the collector thread -> host gather -> task -> H2D -> codelet -> D2H -> slices)
in three admission modes:
  padded   : every request pre-padded to 512 by the client (the reference's fixed-shape contract)
  varlen   : variable_length admission, a batch pads to its longest request (arrival order)
  bucketed : variable_length + length_bucket: the collector batches per ceil(L / bucket)
Host wall clock from the first enqueue to the last completion; one JSON line per mode."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402

from starpu_inference_server_b200 import capi, hostapi, synthetic  # noqa: E402
from starpu_inference_server_b200.bert import BertConfig, BertReplica  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--requests", type=int, default=3840)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--seq", type=int, default=512)
    ap.add_argument("--min-len", type=int, default=16)
    ap.add_argument("--bucket", type=int, default=64)
    ap.add_argument("--coalesce-ms", type=int, default=5)
    args = ap.parse_args()
    spec = BertConfig()
    B, S, H = args.batch, args.seq, spec.hidden
    rep = BertReplica(synthetic.make_weights(spec, 0), max_batch=B, max_seq=S, n_workspaces=4,
                      flags=capi.SPU_FORWARD_AUTO_GRAPH)
    rng = np.random.default_rng(0)
    lens = rng.integers(args.min_len, S + 1, size=args.requests)
    ids_full = rng.integers(1, spec.vocab, size=(args.requests, S), dtype=np.int64)
    for mode in ("padded", "varlen", "bucketed"):
        varlen = mode != "padded"
        srv = hostapi.Server(inputs=[("input_ids", "i64", [1, S]), ("attention_mask", "i64", [1, S])],
                             outputs=[("output0", "f32", [1, S, H])], replicas=[rep], model_kind=hostapi.MODEL_BERT,
                             max_batch_size=B, pool_size=12, nworker_per_cuda=2, cuda_pipeline=4, n_workspaces=4,
                             batch_coalesce_timeout_ms=args.coalesce_ms, variable_length=varlen,
                             length_bucket=args.bucket if mode == "bucketed" else 0)
        reqs = []
        for i, n in enumerate(lens):
            if varlen:
                a = np.ascontiguousarray(ids_full[i:i + 1, :n])
                m = np.ones((1, n), np.int64)
            else:
                a = ids_full[i:i + 1].copy()
                a[0, n:] = 0
                m = np.zeros((1, S), np.int64)
                m[0, :n] = 1
            reqs.append((a, m))

        def run(lo, hi):
            tickets = [srv.enqueue(i, 1, list(reqs[i]), [None]) for i in range(lo, hi)]
            seqs, rows = [], []
            for t in tickets:
                (r,) = srv.wait(t)
                assert r.ok == 1, r.error
                seqs.append(r.batch_seq)
                rows.append(r.batch_rows)
            return seqs, rows

        warm = min(args.requests // 4, 8 * B)
        run(0, warm)
        t0 = time.perf_counter()
        seqs, rows = run(warm, args.requests)
        dt = time.perf_counter() - t0
        n = args.requests - warm
        print(json.dumps({"mode": mode, "inferences_per_s": n / dt, "requests": n, "mean_true_len": float(lens[warm:].mean()),
                          "mean_padded_len_of_the_task": float(np.mean(seqs)) if varlen else float(S),
                          "mean_rows_per_task": float(np.mean(rows)), "tasks": int(srv.stats()["tasks"]),
                          "bucket": args.bucket if mode == "bucketed" else 0, "coalesce_ms": args.coalesce_ms}), flush=True)
        srv.close()
    rep.close()


if __name__ == "__main__":
    main()
