#!/usr/bin/env python
"""GPU-box tool: BASELINE north_star's topology -- ONE process, ONE scheduler, one CUDA worker set and one weight
replica per GPU (`use_cuda: [{device_ids: [0..N-1]}]`, /root/reference/src/core/starpu_setup.cpp:389-432
configure_gpu; src/core/inference_runner.cpp:251-275 clone_model_to_gpus) -- instead of bench.py's N processes.

For N in 1, 2, 4, 8 (up to the visible devices): one spu_server over device_ids [0..N-1], per-request host
buffers -> host gather -> task -> eager scheduler -> a CUDA worker of some device -> codelet -> D2H -> slices.
Prints one JSON line per N: inferences/s end to end, tasks per device, speed-up over N=1."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from starpu_inference_server_b200 import capi, hostapi  # noqa: E402
from starpu_inference_server_b200 import synthetic as syn  # noqa: E402
from starpu_inference_server_b200.bert import BertConfig, BertReplica  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--seq", type=int, default=128)
    ap.add_argument("--steps", type=int, default=400, help="tasks per GPU in the timed region")
    ap.add_argument("--workers", type=int, default=2)
    ap.add_argument("--pipeline", type=int, default=4)
    ap.add_argument("--pool-per-gpu", type=int, default=12)
    ap.add_argument("--max-devices", type=int, default=8)
    args = ap.parse_args()
    n_dev = min(torch.cuda.device_count(), args.max_devices)
    spec = BertConfig()
    B, S = args.batch, args.seq
    w = syn.make_weights(spec, 0)
    flags = capi.SPU_PREC_BF16 | capi.SPU_FORWARD_AUTO_GRAPH
    reps = [BertReplica(w, device=d, max_batch=B, max_seq=S, n_workspaces=args.workers, flags=flags) for d in range(n_dev)]
    host = [syn.make_requests(spec, B, S, seed=1234 + i) for i in range(4)]
    rows = [[(np.ascontiguousarray(i[r:r + 1]), np.ascontiguousarray(m[r:r + 1])) for r in range(B)] for i, m in host]
    io = dict(inputs=[("input_ids", "i64", [1, S]), ("attention_mask", "i64", [1, S])],
              outputs=[("output0", "f32", [1, S, spec.hidden])])
    base = None
    for n in [k for k in (1, 2, 4, 8) if k <= n_dev]:
        srv = hostapi.Server(**io, replicas=reps[:n], device_ids=list(range(n)), model_kind=hostapi.MODEL_BERT,
                             max_batch_size=B, pool_size=args.pool_per_gpu * n, nworker_per_cuda=args.workers,
                             cuda_pipeline=args.pipeline, n_workspaces=args.workers)
        prepared = [srv.prepare_batch([(r, 1, list(a), [None]) for r, a in enumerate(rr)]) for rr in rows]

        def run(n_tasks):
            per_dev, fails = {}, 0
            tickets = [srv.submit_prepared(prepared[i % 4]) for i in range(n_tasks)]
            for t in tickets:
                res = srv.wait(t, cap=B)
                fails += sum(0 if r.ok == 1 else 1 for r in res)
                per_dev[res[0].device_id] = per_dev.get(res[0].device_id, 0) + 1
            return per_dev, fails

        run(16 * n * args.workers)  # first sightings: eager pass + graph capture per (worker, shape)
        segs = []
        for _ in range(3):
            t0 = time.perf_counter()
            per_dev, fails = run(args.steps * n)
            segs.append(time.perf_counter() - t0)
        srv.close()
        dt = sorted(segs)[1]
        rate = B * args.steps * n / dt
        base = base or rate
        print(json.dumps({"devices": n, "topology": "one process, one scheduler, one replica + %d CUDA workers per GPU" % args.workers,
                          "e2e_inferences_per_s": rate, "speedup_vs_1": rate / base, "tasks_per_device": per_dev,
                          "failed": fails, "segments_inf_s": [B * args.steps * n / s for s in segs],
                          "host_cores": len(os.sched_getaffinity(0))}), flush=True)
    for r in reps:
        r.close()


if __name__ == "__main__":
    main()
