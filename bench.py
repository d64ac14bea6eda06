#!/usr/bin/env python
"""Benchmark of the hot path: BERT-base seq128 inferences/s through libspu_b200.so.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch 32] [--seq 128]
    python bench.py --impl reference ...      # the reference's CPU codelet path

A "step" is one pass of the hot path over one batch of synthetic requests: what the
reference's cuda_inference_func does for one StarPU task (src/core/starpu_setup.cpp:
807-846) -- ids/mask [B,S] int64 in, last_hidden_state [B,S,768] fp32 out.
Workload = BASELINE.json configs[1]: BERT-base, seq 128, dynamic batch at its cap 32,
bf16, one B200 CUDA worker per GPU.  N>1: one process per GPU (torchrun), one weight
replica each, requests are independent -> weak scaling, no collective on the data path
(torch.distributed is used only for the timing barrier / max-over-ranks).

Prints ONE JSON line (rank 0).  value = whole-job inferences/s with inputs resident in
HBM; e2e = same metric through the host-buffer API (pinned host ids/mask -> H2D ->
forward -> D2H of output0) with copies inside the timed region.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

METRIC = "BERT-base seq128 inferences/s"
UNIT = "inferences/s"
LAYERS, HIDDEN, FFN = 12, 768, 3072


def gemm_flops_per_seq(seq):
    return LAYERS * 24 * seq * HIDDEN * HIDDEN  # Lyr*24*L*H^2


def attn_flops_per_seq(seq):
    return LAYERS * 4 * seq * seq * HIDDEN


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"tflops": float(d.get("bf16_tflops_sustained", d.get("bf16_tflops"))),
                "hbm_gbs": float(d["hbm_gbs"]), "source": "measured (MEASURED_PEAKS.json, sustained bf16)"}
    return {"tflops": 1400.0, "hbm_gbs": 6650.0, "source": "fallback (B200_PROFILING.md)"}


class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._halt = threading.Event()
        self.ok = False
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
                 nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: "hw_power_brake"}
        while not self._halt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.02)

    def stop(self):
        self._halt.set()
        if self.ok:
            self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def make_batches(cfg, batch, seq, n_sets, seed0=1234):
    from starpu_inference_server_b200 import synthetic

    return [synthetic.make_requests(cfg, batch, seq, seed=seed0 + i) for i in range(n_sets)]


# ----------------------------------------------------------------- reference arm
def cpu_codelet_forward(batch, seq, budget_s, max_steps, warmup, threads=None):
    """The reference's CPU codelet path restated: TorchScript module forward on the host
    cores + copy of the output into the (host) StarPU buffer
    (src/core/starpu_setup.cpp:784-801, tensor_builder.cpp:161-189).  The TorchScript
    artifact is the traced oracle module (HF BertModel does not trace under
    transformers 5.5).  Returns (inferences/s, ms list, steps, threads)."""
    from oracle import bert_oracle as bo

    threads = threads or os.cpu_count() or 1
    torch.set_num_threads(threads)
    spec = bo.BERT_BASE
    w = bo.make_weights(spec, 0)
    ids, mask = bo.make_inputs(spec, batch, seq, seed=1234)
    ids_t, mask_t = torch.from_numpy(ids), torch.from_numpy(mask)
    m = bo.BertOracleModule(w, spec).eval()
    with torch.no_grad():
        ts = torch.jit.freeze(torch.jit.trace(m, (ids_t, mask_t)))
    out_buf = torch.empty(batch, seq, spec.hidden)
    times = []
    with torch.inference_mode():
        for _ in range(max(1, warmup)):
            out_buf.copy_(ts(ids_t, mask_t))
        t_begin = time.perf_counter()
        for _ in range(max_steps):
            t0 = time.perf_counter()
            out_buf.copy_(ts(ids_t, mask_t))
            times.append((time.perf_counter() - t0) * 1e3)
            if time.perf_counter() - t_begin > budget_s:
                break
    total_s = sum(times) / 1e3
    return batch * len(times) / total_s, times, len(times), threads


def run_reference(args, rank, world):
    if rank != 0:
        return
    steps = args.steps
    value, times, done, threads = cpu_codelet_forward(args.batch, args.seq, budget_s=150.0,
                                                      max_steps=steps, warmup=min(args.warmup, 2))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": done, "warmup": min(args.warmup, 2), "ms_per_step": float(np.mean(times)),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": f"BERT-base seq{args.seq} batch {args.batch} (BASELINE configs[1]) on the host cores",
                   "batch": args.batch, "seq_len": args.seq,
                   "path": "TorchScript forward of the traced oracle module (CPU codelet restatement)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"{done} forward passes of batch {args.batch} x seq {args.seq}"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "latency_ms": {"p50": float(np.percentile(times, 50)), "p99": float(np.percentile(times, 99))},
    }
    print(json.dumps(line), flush=True)


def reduce_max_over_ranks(values, device):
    """Slowest rank's time for each entry (the contract's max-over-ranks); identity at N=1.
    Ranks are independent replicas: this all-reduce of a few scalars is the only collective
    of the whole benchmark and it is outside every timed region."""
    import torch.distributed as dist

    t = torch.tensor(list(values), device=device, dtype=torch.float64)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(v) for v in t]


def whole_job_rate(units_per_rank_step, steps, world, slowest_ms):
    """Weak scaling: every rank processes its own `units_per_rank_step` x steps; the job's
    throughput is all ranks' units over the slowest rank's time."""
    return world * units_per_rank_step * steps / (slowest_ms / 1e3)


# ---------------------------------------------------------------------- B200 arm
def run_b200(args, rank, local_rank, world):
    import torch.distributed as dist

    from starpu_inference_server_b200 import capi, synthetic
    from starpu_inference_server_b200.bert import BertConfig, BertReplica

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the hot path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        # replicas only: the ranks never exchange data, so the rendezvous is a
        # host-side gloo group used for the barrier and the max-over-ranks of a few scalars;
        # NCCL is deliberately not initialised (and cannot pollute stdout with its banner)
        dist.init_process_group("gloo")
    assert capi.lib().spu_device_is_sm100(local_rank) == 1, "device is not sm_100"

    spec = BertConfig()
    B, S = args.batch, args.seq
    n_pipe = 4  # workspaces / CUDA workers, like STARPU_NWORKER_PER_CUDA=4 (models/bert.yml:6-8)
    flags = {"bf16": capi.SPU_PREC_BF16, "fp8": capi.SPU_PREC_FP8_E4M3,
             "fp8x6": capi.SPU_PREC_FP8_E4M3 | (6 << 8)}[args.precision]
    if args.output_bf16:
        flags |= capi.SPU_OUTPUT_BF16
    if args.auto_graph:
        # the codelet path replays a CUDA graph from the second sighting of a slot's device
        # buffers on (warm-up capture, src/core/warmup.cpp:225-299)
        flags |= capi.SPU_FORWARD_AUTO_GRAPH
    w = synthetic.make_weights(spec, 0)
    rep = BertReplica(w, device=local_rank, max_batch=B, max_seq=S, n_workspaces=n_pipe, flags=flags)
    del w
    n_sets = 4
    from starpu_inference_server_b200 import synthetic as syn

    host = [syn.make_requests(spec, B, S, seed=1234 + 100 * rank + i, min_len=args.min_len) for i in range(n_sets)]
    true_tokens = float(np.mean([m.sum() for _, m in host]))  # FLOPs are charged on unpadded tokens
    ids_d = [torch.from_numpy(i).to(dev) for i, _ in host]
    mask_d = [torch.from_numpy(m).to(dev) for _, m in host]
    out_d = [torch.empty(B, S, spec.hidden, device=dev, dtype=torch.bfloat16 if args.output_bf16 else torch.float32)
             for _ in range(n_pipe)]
    stream = torch.cuda.Stream(device=dev)
    if flags & capi.SPU_PREC_FP8_E4M3:
        # per-tensor activation scales from a 32-sequence calibration draw
        ci, cm = syn.make_requests(spec, min(32, B), S, seed=999, min_len=args.min_len)
        rep.calibrate_fp8(torch.from_numpy(ci).to(dev), torch.from_numpy(cm).to(dev), stream=stream)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---------------- device-resident throughput (value) ----------------
    # warm-up hook (src/core/warmup.cpp:225-299): one CUDA graph per (worker, input set)
    # pointer tuple; `--streams` workers issue tasks round-robin like StarPU's CUDA workers
    n_str = max(1, min(args.streams, n_pipe))
    wstreams = [stream] + [torch.cuda.Stream(device=dev) for _ in range(n_str - 1)]
    for k, st in enumerate(wstreams):
        with torch.cuda.stream(st):
            for i in range(n_sets):
                rep.capture(ids_d[i], mask_d[i], out_d[k], ws=k, stream=st)
    torch.cuda.synchronize(dev)

    def step_resident(i, k=0):
        rep.forward_async(ids_d[i % n_sets], mask_d[i % n_sets], out_d[k], ws=k, stream=wstreams[k])

    for i in range(args.warmup):
        step_resident(i, i % n_str)
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for st in wstreams[1:]:
        st.wait_event(e0)
    for i in range(args.steps):
        step_resident(i, i % n_str)
    for st in wstreams[1:]:
        ev = torch.cuda.Event()
        ev.record(st)
        stream.wait_event(ev)
    e1.record(stream)
    barrier()
    clocks = sampler.stop()
    ms_total = e0.elapsed_time(e1)

    # ---------------- per-task device latency (one batch, idle GPU) ----------
    lat = []
    with torch.cuda.stream(stream):
        for i in range(min(args.steps, 200)):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            step_resident(i)
            b.record(stream)
            b.synchronize()
            lat.append(a.elapsed_time(b))

    # ---------------- per-kernel-class device time (roofline) ----------------
    # each class runs alone as 5 forwards' worth of launches inside one CUDA graph
    stage_ms = {k: min(rep.stage_chain_ms(ids_d[0], mask_d[0], out_d[0], [k], stream=stream) for _ in range(3))
                for k in capi.STAGE_NAMES}
    gemm_names = ["gemm_qkv", "gemm_attn_out", "gemm_ffn1", "gemm_ffn2"]
    gemm_ms = min(rep.stage_chain_ms(ids_d[0], mask_d[0], out_d[0], gemm_names, stream=stream) for _ in range(3))
    launches = rep.launches_per_forward

    # ---------------- end-to-end through the codelet / task path (e2e) ----------------
    # per-request host buffers ([1,S] ids + mask each) -> host gather into a pinned slot ->
    # StarPU-shaped task -> H2D -> codelet cuda_func -> C-ABI kernels -> D2H into the pinned
    # output slot -> per-request slices; everything inside the timed region.
    from starpu_inference_server_b200 import hostapi

    # pool 12 as models/bert.yml; 2 CUDA workers x STARPU_CUDA_PIPELINE 4 (the reference's pipeline depth) measured
    # best of the swept (workers, pipeline) pairs at N=1 and N=8 (profiles/r06_e2e_config.md); env knobs: sweeps only
    e2e_workers = min(n_pipe, int(os.environ.get("SPU_BENCH_WORKERS", 2)))
    e2e_pipeline = int(os.environ.get("SPU_BENCH_PIPELINE", 4))
    e2e_pool = int(os.environ.get("SPU_BENCH_POOL", 12))
    io = dict(inputs=[("input_ids", "i64", [1, S]), ("attention_mask", "i64", [1, S])],
              outputs=[("output0", "bf16" if args.output_bf16 else "f32", [1, S, spec.hidden])])
    srv = hostapi.Server(**io, replicas=[rep], device_ids=[local_rank], model_kind=hostapi.MODEL_BERT,
                         max_batch_size=B, pool_size=e2e_pool, nworker_per_cuda=e2e_workers, cuda_pipeline=e2e_pipeline,
                         n_workspaces=n_pipe)
    req_rows = [[(np.ascontiguousarray(i[r:r + 1]), np.ascontiguousarray(m[r:r + 1])) for r in range(B)]
                for i, m in host]
    prepared = [srv.prepare_batch([(r, 1, [a, b], [None]) for r, (a, b) in enumerate(rows)]) for rows in req_rows]
    checksum, task_ms, fails = 0.0, [], 0

    def run_e2e(n_steps, record):
        nonlocal checksum, fails
        tickets = [srv.submit_prepared(prepared[i % n_sets]) for i in range(n_steps)]
        for t in tickets:
            for r in srv.wait(t, cap=B):
                fails += 0 if r.ok == 1 else 1
                if record:
                    checksum += float(r.first_value)
                    task_ms.append(r.task_ms)

    # auto-graph: every (worker, slot) tuple has to be seen twice before it replays a graph
    run_e2e(max(args.warmup, 6 * e2e_workers * e2e_pool if args.auto_graph else 20), False)
    # three back-to-back segments of K steps each, the MEDIAN segment is reported: one host-side hiccup
    # (a descheduled worker thread) inside a 0.2 s region otherwise moves the number by 20 %
    e2e_segments = []
    for _ in range(3):
        barrier()
        t0 = time.perf_counter()
        run_e2e(args.steps, True)
        e2e_segments.append(time.perf_counter() - t0)
        barrier()
    e2e_s = sorted(e2e_segments)[1]
    host_stats = srv.stats()
    srv.close()
    rep.close()
    if fails:
        raise SystemExit(f"bench.py: {fails} requests failed on the e2e path")

    # ---------------- reduce over ranks ----------------
    ms_total, e2e_ms = reduce_max_over_ranks([ms_total, e2e_s * 1e3], torch.device("cpu"))
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peaks = load_peaks()
    if args.precision == "fp8":
        # no measured FP8 figure in MEASURED_PEAKS.json: 2x the measured bf16 one (the nominal FP8:bf16 ratio)
        peaks["tflops"] *= 2.0
        peaks["source"] += " x 2 for FP8 E4M3 (nominal ratio, not measured)"
    value = whole_job_rate(B, args.steps, world, ms_total)
    e2e_value = whole_job_rate(B, args.steps, world, e2e_ms)
    gemm_flop_step = LAYERS * 24 * true_tokens * HIDDEN * HIDDEN  # charged on true lengths
    gemm_tflops = gemm_flop_step / (gemm_ms / 1e3) / 1e12
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16" if args.precision == "bf16" else "fp8_e4m3",
        "data": "synthetic",
        "config": {"workload": f"BERT-base seq{S} dynamic batch at cap {B}, {args.precision}, one weight replica per B200 "
                               + ("(BASELINE configs[1])" if (B, S, args.precision, args.min_len) == (32, 128, "bf16", None)
                                  else "(variant of BASELINE configs[1])")
                               + "; random-init weights seed 0, ids U[0,30522), "
                               + ("all-ones mask" if args.min_len is None else f"lengths U[{args.min_len},{S}] padded+masked"),
                   "batch": B, "seq_len": S, "replicas": world, "output_dtype": "bf16" if args.output_bf16 else "f32", "resident_streams": n_str,
                   "mean_true_tokens_per_batch": true_tokens,
                   "l2": "no explicit flush: each step streams 170 MB of bf16 weights plus ~110 MB of activations "
                         "(> 126 MB L2) and input batches rotate over 4 distinct buffers"},
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 2 * B * S * 8,
                "d2h_bytes_per_step": B * S * HIDDEN * (2 if args.output_bf16 else 4), "checksum": checksum,
                "path": "libspu_host.so: per-request host buffers -> host gather -> task -> H2D -> codelet -> D2H -> "
                        f"slices ({e2e_workers} CUDA workers x pipeline {e2e_pipeline}, pool {e2e_pool}, "
                        + ("CUDA-graph replay per (worker, slot) tuple)" if args.auto_graph else "eager launches)"),
                "tasks": int(host_stats["tasks"]),
                "how": f"median of 3 segments of {args.steps} steps (rank 0 segments, inferences/s: "
                       + ", ".join(f"{B * args.steps * world / t:.0f}" for t in e2e_segments) + ")",
                "task_latency_ms": {"p50": float(np.percentile(task_ms, 50)), "p99": float(np.percentile(task_ms, 99)),
                                    "what": "submit -> output callback per task, queueing included"}},
        "gpu_launches": launches * args.steps,
        "latency_ms": {"p50": float(np.percentile(lat, 50)), "p99": float(np.percentile(lat, 99)),
                       "what": "device time of one task (batch) = codelet_start..codelet_end"},
        "roofline": {"bound": "tensor", "kernel": "gemm_tcgen05_kernel (QKV, attn-out, FFN1, FFN2; 48 launches/step)",
                     "achieved": gemm_tflops, "peak": peaks["tflops"], "unit": "TFLOP/s",
                     "frac": gemm_tflops / peaks["tflops"], "peak_source": peaks["source"], **gemm_traffic(B, S, args.precision),
                     "flops_per_step": gemm_flop_step, "kernel_ms_per_step": gemm_ms,
                     "how": "48 GEMM launches x 5 forwards back to back in one CUDA graph, CUDA events on the launch stream"},
        "stage_ms_per_step": stage_ms,
    }
    if world == 1 and not args.no_cpu_baseline:
        v, times, done, threads = cpu_codelet_forward(B, S, budget_s=20.0, max_steps=20, warmup=1)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                                "sample": f"{done} forward passes of batch {B} x seq {S} "
                                          "(TorchScript fp32 on the host cores)"}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def gemm_traffic(B, S, precision):
    """roofline.traffic: dram__bytes_read+write per GEMM launch (mean of the four GEMMs of a layer) from the committed
    `ncu --set full` capture, when this run is the workload the capture was taken on; else null."""
    p = os.path.join(ROOT, "profiles", "gemm_traffic.json")
    try:
        t = json.load(open(p))
        w = t["workload"]
        if (w["batch"], w["seq_len"], w["precision"]) == (B, S, precision):
            return {"traffic": t["mean_dram_bytes_per_launch"], "traffic_unit": "bytes per GEMM launch (ncu dram read+write, cold)",
                    "traffic_source": t["source"]}
    except (OSError, KeyError, ValueError):
        pass
    return {"traffic": None}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--seq", type=int, default=128)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp8", "fp8x6"])
    ap.add_argument("--min-len", type=int, default=None, help="variable-length requests: lengths U[min_len, seq]")
    ap.add_argument("--streams", type=int, default=2, help="CUDA workers issuing resident tasks round-robin")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--output-bf16", action="store_true",
                    help="variant: output0 as bf16 (SPU_OUTPUT_BF16, SURVEY 8(f) row 3): half the response bytes")
    ap.add_argument("--no-auto-graph", dest="auto_graph", action="store_false",
                    help="e2e path launches every kernel eagerly instead of replaying CUDA graphs captured on the second "
                         "sighting of a (worker, slot) tuple (SPU_FORWARD_AUTO_GRAPH; profiles/r06_e2e_config.md: "
                         "33.7k vs 32.2k inf/s at N=1, 259k vs 220k at N=8 where host cores are scarce)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        if args.steps > 40:
            args.steps = 40  # bounded sample: ~1 s per batch-32 step on 8 cores
        run_reference(args, rank, world)
    else:
        run_b200(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
