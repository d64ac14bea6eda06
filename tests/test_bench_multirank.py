"""CPU, world_size 2 over gloo: the N>1 logic of bench.py -- ranks are independent replicas,
the only collective is the max-over-ranks of the timed intervals, value = all ranks' units
over the slowest rank's time, and the reference arm prints on rank 0 only."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r"""
import json, os, sys
sys.path.insert(0, %r)
import torch, torch.distributed as dist
import bench
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
# rank r "measured" (10 + 5 r) ms for the resident loop and (20 - 3 r) ms end to end
vals = bench.reduce_max_over_ranks([10.0 + 5 * rank, 20.0 - 3 * rank], torch.device("cpu"))
rate = bench.whole_job_rate(32, 100, world, vals[0])
# one file per rank: two ranks writing to the same stdout pipe can interleave mid-line
with open(os.path.join(os.environ["SPU_TEST_OUT"], "rank" + str(rank) + ".json"), "w") as f:
    json.dump({"rank": rank, "vals": vals, "rate": rate}, f)
dist.destroy_process_group()
""" % ROOT


def test_max_over_ranks_and_weak_scaling_rate(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", SPU_TEST_OUT=str(tmp_path))
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29731", str(script)],
                         capture_output=True, text=True, env=env, timeout=240)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [json.loads((tmp_path / f"rank{r}.json").read_text()) for r in range(2)]
    assert sorted(l["rank"] for l in lines) == [0, 1]
    for l in lines:
        assert l["vals"] == [15.0, 20.0]                      # slowest rank of each interval
        assert l["rate"] == pytest.approx(2 * 32 * 100 / 0.015)  # both replicas' units / slowest time


def test_single_rank_is_identity():
    import torch

    sys.path.insert(0, ROOT)
    import bench

    assert bench.reduce_max_over_ranks([1.5, 2.5], torch.device("cpu")) == [1.5, 2.5]
    assert bench.whole_job_rate(32, 10, 1, 1000.0) == 320.0


def test_reference_arm_other_ranks_exit_quietly():
    """bench.py --impl reference under torchrun: ranks != 0 print nothing and exit 0."""
    env = dict(os.environ, RANK="1", LOCAL_RANK="1", WORLD_SIZE="2")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2",
                          "--steps", "1", "--warmup", "1"], capture_output=True, text=True, env=env, timeout=240)
    assert out.returncode == 0 and out.stdout.strip() == ""


def test_config_objects_are_identical_in_both_arms_and_name_the_baseline_workload():
    sys.path.insert(0, ROOT)
    import argparse

    import bench

    for cid in "ABCDE":
        args = argparse.Namespace(config=cid, batch=None, seq=None, precision=None, min_len=None)
        cfg = bench.resolve_config(args)
        a, b = bench.workload_config(cfg, 8), bench.workload_config(dict(cfg), 8)
        assert a == b and a["config_id"] == cid and "BASELINE configs[" in a["workload"] and not cfg["variant"]
        assert set(a) == {"workload", "config_id", "model", "batch", "seq_len", "precision", "min_len", "replicas", "l2"}
    v = bench.resolve_config(argparse.Namespace(config="B", batch=64, seq=None, precision=None, min_len=None))
    assert v["variant"] and "variant" in bench.workload_config(v, 1)["workload"]
    assert bench.metric_name(bench.resolve_config(argparse.Namespace(config="B", batch=None, seq=None, precision=None,
                                                                      min_len=None))) == "BERT-base seq128 inferences/s"


def test_parity_block_counts_what_it_says():
    import numpy as np

    sys.path.insert(0, ROOT)
    import bench

    rng = np.random.default_rng(0)
    ref = rng.standard_normal((4, 8, 16)).astype(np.float32)
    got = ref.copy()
    got[0, 0, 0] += 0.5          # one element off: breaks the client rule once, may flip no argmax
    got[1, 1] = -got[1, 1]       # one row whose argmax flips
    live = np.ones((4, 8), dtype=np.int64)
    live[3, 4:] = 0              # padded rows are not counted
    got[3, 5] = 100.0
    p = bench.parity_block(got, ref, "bf16", live)
    assert p["rows"] == 28 and p["client_rule"]["violations"] >= 17
    assert p["max_abs_over_max_ref"] == np.abs(got - ref)[live.astype(bool)].max() / np.abs(ref[live.astype(bool)]).max()
    assert p["top1_strict"] == 27 / 28 and not p["gate_max_rel_ok"]
    assert bench.parity_block(ref, ref, "fp8")["gate_max_rel_ok"] and bench.parity_block(ref, ref, "fp8")["top1_strict"] == 1.0
