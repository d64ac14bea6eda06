#!/usr/bin/env python
"""A/B of two checkouts of this repo on ONE box (was round 2 slower than round 1, or was the box?).

    python tools/sweeps/r3_ab_chain.py <checkout root> <label>

Loads the package (and its in-tree libspu_b200.so) from <checkout root>, builds a BERT-base replica from seeded
weights and prints one JSON line: device ms per forward of every GEMM class alone, of the GEMM chain, and of the
whole forward at batch 32 x seq 128 (spu_bert_stage_chain_ms: 5 forwards' worth of launches in one CUDA graph, CUDA
events on the launch stream, best of 5 after 1 s of idle).  One process per checkout: the two libraries export the
same symbols.  tools/sweeps/r3_bisect.sh interleaves the two.
"""
import json
import sys
import time

root, label = sys.argv[1], sys.argv[2]
sys.path.insert(0, root)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from starpu_inference_server_b200 import capi  # noqa: E402
from starpu_inference_server_b200.bert import BertConfig, BertReplica  # noqa: E402
from starpu_inference_server_b200 import synthetic as syn  # noqa: E402

cfg = BertConfig()
state = syn.make_state(cfg, seed=0) if hasattr(syn, "make_state") else syn.make_weights(cfg, seed=0)
rep = BertReplica(state, cfg, max_batch=32, max_seq=128, n_workspaces=1)
rng = np.random.default_rng(1)
ids = torch.from_numpy(rng.integers(0, cfg.vocab, size=(32, 128), dtype=np.int64)).cuda()
mask = torch.ones_like(ids)
out = torch.empty((32, 128, cfg.hidden), device="cuda")
stream = torch.cuda.Stream()
gemms = ["gemm_qkv", "gemm_attn_out", "gemm_ffn1", "gemm_ffn2"]
with torch.cuda.stream(stream):
    for _ in range(3):
        rep.forward_async(ids, mask, out, stream=stream)
torch.cuda.synchronize()
res = {"label": label, "lib": capi.lib()._name}
for rnd in range(2):
    time.sleep(1.0)
    res[f"chain_ms_{rnd}"] = min(rep.stage_chain_ms(ids, mask, out, gemms, stream=stream) for _ in range(5))
    for g in gemms:
        res[f"{g}_us_{rnd}"] = min(rep.stage_chain_ms(ids, mask, out, [g], stream=stream) for _ in range(5)) / 12 * 1e3
    res[f"forward_ms_{rnd}"] = min(rep.stage_chain_ms(ids, mask, out, None, stream=stream) for _ in range(5))
print(json.dumps(res), flush=True)
