#!/usr/bin/env python
"""Target for `ncu`: two eager BERT-base forwards (batch 32 x seq 128 by default) through the
C-ABI.  Skip the first forward's launches and capture the second's.  `packed MIN_LEN`: right-padded
batch with lengths U[MIN_LEN, S] through the packed forward (BASELINE configs[2])."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from starpu_inference_server_b200 import synthetic  # noqa: E402
from starpu_inference_server_b200.bert import BertConfig, BertReplica  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
S = int(sys.argv[2]) if len(sys.argv) > 2 else 128
packed = len(sys.argv) > 3 and sys.argv[3] == "packed"
min_len = int(sys.argv[4]) if len(sys.argv) > 4 else None
spec = BertConfig()
rep = BertReplica(synthetic.make_weights(spec, 0), max_batch=B, max_seq=S)
ids, mask = synthetic.make_requests(spec, B, S, seed=1, min_len=min_len)
lens = mask.sum(1)
ids, mask = torch.from_numpy(ids).cuda(), torch.from_numpy(mask).cuda()
out = torch.empty(B, S, spec.hidden, device="cuda")
for _ in range(2):
    if packed:
        rep.forward_packed_async(ids, lens, out)
    else:
        rep.forward_async(ids, mask, out)
    torch.cuda.synchronize()
rep.close()
