#!/usr/bin/env python
"""GPU-box tool: ViT-L/16 224x224 images/s through the C-ABI (BASELINE configs[4]): resident
device loop (CUDA events) and the end-to-end codelet/task path with host pixel buffers."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from starpu_inference_server_b200 import hostapi  # noqa: E402
from starpu_inference_server_b200.vit import VitConfig, VitReplica  # noqa: E402


def make_weights(cfg, seed=0):
    """Same stream as oracle.vit_oracle.make_weights (kept local: tools never import the oracle)."""
    import math

    from starpu_inference_server_b200.vit import _tensor_order

    rng = np.random.Generator(np.random.PCG64(seed))
    H, M, S = cfg.hidden, cfg.mlp, cfg.seq
    shapes = {"conv_proj.weight": (H, cfg.channels, cfg.patch, cfg.patch), "conv_proj.bias": (H,),
              "class_token": (1, 1, H), "encoder.pos_embedding": (1, S, H), "encoder.ln.weight": (H,),
              "encoder.ln.bias": (H,), "heads.head.weight": (cfg.classes, H), "heads.head.bias": (cfg.classes,)}
    out = {}
    for name in _tensor_order(cfg.layers):
        if name in shapes:
            shape = shapes[name]
        elif name.endswith("in_proj_weight"):
            shape = (3 * H, H)
        elif name.endswith("in_proj_bias"):
            shape = (3 * H,)
        elif name.endswith("mlp.0.weight"):
            shape = (M, H)
        elif name.endswith("mlp.0.bias"):
            shape = (M,)
        elif name.endswith("mlp.3.weight"):
            shape = (H, M)
        elif name.endswith("out_proj.weight"):
            shape = (H, H)
        else:
            shape = (H,)
        a = rng.standard_normal(shape, dtype=np.float32)
        is_ln = ".ln" in name or "ln_" in name
        if name.endswith("weight") and is_ln:
            a = np.float32(1.0) + a * np.float32(0.02)
        elif name.endswith("bias"):
            a *= np.float32(0.02)
        elif name.startswith("heads."):
            a *= np.float32(0.05)
        elif a.ndim == 2 or name == "conv_proj.weight":
            a *= np.float32(0.5 / math.sqrt(int(np.prod(shape[1:]))))
        else:
            a *= np.float32(0.02)
        out[name] = a
    return out


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 30
    # under torchrun: one replica + process per GPU, no collective on the data path (gloo barrier + max of times)
    import torch.distributed as dist
    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", 1), ("RANK", 0), ("LOCAL_RANK", 0)))
    if world > 1:
        dist.init_process_group("gloo")
    torch.cuda.set_device(local)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def slowest(v):
        t = torch.tensor([v], dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    cfg = VitConfig()
    rep = VitReplica(make_weights(cfg), cfg, device=local, max_batch=B, n_workspaces=4)
    rng = np.random.Generator(np.random.PCG64(7))
    host = [rng.random((B, 3, 224, 224), dtype=np.float32) for _ in range(2)]
    x = [torch.from_numpy(h).cuda() for h in host]
    out = torch.empty(B, cfg.classes, device="cuda")
    s = torch.cuda.Stream()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(s):
        for i in range(3):
            rep.forward_async(x[i % 2], out, stream=s)
    barrier()
    with torch.cuda.stream(s):
        e0.record(s)
        for i in range(steps):
            rep.forward_async(x[i % 2], out, stream=s)
        e1.record(s)
    barrier()
    ms = slowest(e0.elapsed_time(e1)) / steps
    gemm_flop = (2 * 196 * 768 * 1024 + 24 * 24 * 197 * 1024 * 1024 + 2 * 1024 * 1000) * B
    res = {"workload": f"ViT-L/16 224x224 batch {B} bf16 (BASELINE configs[4])", "ms_per_step": ms,
           "images_per_s": world * B / (ms / 1e3), "n_gpus": world,
           "dense_tflops_whole_forward_per_gpu": gemm_flop / (ms / 1e3) / 1e12,
           "launches_per_forward": rep.launches_per_forward}
    # end to end: [1,3,224,224] host requests -> gather -> task -> H2D -> codelet -> D2H -> slices
    srv = hostapi.Server(**hostapi.VIT_L_IO, replicas=[rep], device_ids=[local], model_kind=hostapi.MODEL_VIT, max_batch_size=B,
                         pool_size=6, nworker_per_cuda=4, cuda_pipeline=2, n_workspaces=4)
    rows = [[np.ascontiguousarray(h[r:r + 1]) for r in range(B)] for h in host]
    prepared = [srv.prepare_batch([(r, 1, [a], [None]) for r, a in enumerate(rr)]) for rr in rows]
    for t in [srv.submit_prepared(prepared[i % 2]) for i in range(4)]:
        srv.wait(t, cap=B)
    barrier()
    t0 = time.perf_counter()
    tickets = [srv.submit_prepared(prepared[i % 2]) for i in range(steps)]
    ok = all(r.ok == 1 for t in tickets for r in srv.wait(t, cap=B))
    dt = slowest(time.perf_counter() - t0)
    res["e2e_images_per_s"] = world * B * steps / dt
    res["e2e_ok"] = ok
    res["h2d_bytes_per_step"] = B * 3 * 224 * 224 * 4
    srv.close()
    rep.close()
    if rank == 0:
        print(json.dumps(res))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
