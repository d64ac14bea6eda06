#!/usr/bin/env python
"""GPU-box tool: host cost of enqueueing one BERT-base forward eagerly (85 launches, tensor maps from the plan's
cache) against replaying its CUDA graph, and the device latency of both, at b=1 and b=32."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from starpu_inference_server_b200 import synthetic as syn  # noqa: E402
from starpu_inference_server_b200.bert import BertConfig, BertReplica  # noqa: E402


def main():
    spec = BertConfig()
    rep = BertReplica(syn.make_weights(spec, 0), max_batch=32, max_seq=128, n_workspaces=2)
    s = torch.cuda.Stream()
    for B in (1, 8, 32):
        ids, mask = syn.make_requests(spec, B, 128, seed=3)
        ids_d, mask_d = torch.from_numpy(ids).cuda(), torch.from_numpy(mask).cuda()
        out = torch.empty(B, 128, spec.hidden, device="cuda")
        res = {"batch": B}
        for mode, ws in (("eager", 0), ("graph", 1)):
            if mode == "graph":
                rep.capture(ids_d, mask_d, out, ws=ws, stream=s)
            for _ in range(5):
                rep.forward_async(ids_d, mask_d, out, ws=ws, stream=s)
            torch.cuda.synchronize()
            host, dev = [], []
            for _ in range(50):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(s)
                t0 = time.perf_counter()
                rep.forward_async(ids_d, mask_d, out, ws=ws, stream=s)
                host.append((time.perf_counter() - t0) * 1e6)
                e1.record(s)
                e1.synchronize()
                dev.append(e0.elapsed_time(e1) * 1e3)
            res[mode] = {"host_enqueue_us_p50": float(np.percentile(host, 50)), "device_us_p50": float(np.percentile(dev, 50))}
        print(json.dumps(res), flush=True)
    rep.close()


if __name__ == "__main__":
    main()
