#!/usr/bin/env python
"""Target for `ncu` (tools/sweeps/r3_ncu.sh): the device-resident gather / scatter kernels at the benchmark shape.
32 requests x 128 tokens in DEVICE memory through spu_bert_forward_requests_checked_async (gather_pad_mask_kernel,
scatter_requests_kernel) and the same rows through spu_gather_segments_async / spu_scatter_segments_async
(segcopy_kernel), twice each; the second pass is the one to read."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from starpu_inference_server_b200 import capi, synthetic  # noqa: E402
from starpu_inference_server_b200.bert import BertConfig, BertReplica  # noqa: E402

B, S = 32, 128
spec = BertConfig()
rep = BertReplica(synthetic.make_weights(spec, 0), max_batch=B, max_seq=S, flags=capi.SPU_REQUEST_IO)
rng = np.random.default_rng(0)
ids = [torch.from_numpy(rng.integers(0, spec.vocab, size=S).astype(np.int64)).cuda() for _ in range(B)]
mask = [torch.ones(S, dtype=torch.int64, device="cuda") for _ in range(B)]
outs = [torch.empty(S, spec.hidden, device="cuda") for _ in range(B)]
table = lambda ts: torch.tensor([t.data_ptr() for t in ts], dtype=torch.int64, device="cuda")  # noqa: E731
lens = torch.full((B,), S, dtype=torch.int32, device="cuda")
status = torch.zeros(1, dtype=torch.int32, device="cuda")
it, mt, ot = table(ids), table(mask), table(outs)
for _ in range(2):
    rep.forward_requests_async(it, mt, lens, B, S, ot, status=status)
    torch.cuda.synchronize()
# dense byte gather / scatter of the fp32 output rows (segcopy_kernel): 32 segments of 128 x 768 floats
L = capi.lib()
seg = S * spec.hidden * 4
slot = torch.empty(B * S * spec.hidden, device="cuda")
off = torch.arange(B, dtype=torch.int64, device="cuda") * seg
nb = torch.full((B,), seg, dtype=torch.int64, device="cuda")
st = capi.stream_ptr(None)
for _ in range(2):
    capi.check(L.spu_gather_segments_async(capi.ptr(ot), capi.ptr(off), capi.ptr(nb), B, capi.ptr(slot), st))
    capi.check(L.spu_scatter_segments_async(capi.ptr(slot), capi.ptr(off), capi.ptr(nb), capi.ptr(ot), B, st))
    torch.cuda.synchronize()
assert int(status.item()) == 0
rep.close()
