#!/usr/bin/env python
"""GPU-box diagnostic: error of the CUDA forward against the fp32 oracle, split by valid /
padded tokens, next to the error a bf16-operand emulation of the same graph makes on the
CPU (the floor any bf16 pipeline has).  Not a test; prints JSON lines."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from oracle import bert_oracle as bo  # noqa: E402
from starpu_inference_server_b200.bert import BertReplica  # noqa: E402


def bf(x):
    return x.to(torch.bfloat16).to(torch.float32)


def emulated(w, spec, ids, mask):
    """bert_forward with every GEMM operand rounded to bf16 (fp32 accumulate)."""
    t = {k: torch.from_numpy(v) for k, v in w.items()}
    ids_t, m = torch.from_numpy(ids), torch.from_numpy(mask)
    B, S = ids.shape
    H, nh = spec.hidden, spec.heads
    ln = lambda x, g, b: torch.nn.functional.layer_norm(x, (H,), g, b, spec.ln_eps)
    x = t["embeddings.word_embeddings.weight"][ids_t] + t["embeddings.position_embeddings.weight"][:S] \
        + t["embeddings.token_type_embeddings.weight"][0]
    x = ln(x, t["embeddings.LayerNorm.weight"], t["embeddings.LayerNorm.bias"])
    bias = torch.where(m != 0, 0.0, -1e30)[:, None, None, :]
    for l in range(spec.layers):
        p = f"encoder.layer.{l}."
        lin = lambda a, n: bf(a) @ bf(t[p + n + ".weight"]).T + t[p + n + ".bias"]
        q, k, v = (bf(lin(x, "attention.self." + n)).view(B, S, nh, 64).transpose(1, 2)
                   for n in ("query", "key", "value"))
        pr = torch.softmax(q @ k.transpose(-1, -2) / 8.0 + bias, -1)
        ctx = bf((bf(pr) @ v).transpose(1, 2).reshape(B, S, H))
        x = ln(lin(ctx, "attention.output.dense") + x, t[p + "attention.output.LayerNorm.weight"],
               t[p + "attention.output.LayerNorm.bias"])
        h = bf(torch.nn.functional.gelu(lin(x, "intermediate.dense")))
        x = ln(lin(h, "output.dense") + x, t[p + "output.LayerNorm.weight"], t[p + "output.LayerNorm.bias"])
    return x.numpy()


def stats(name, got, ref, mask):
    out = {"case": name}
    for lab, sel in (("all", np.ones_like(mask, bool)), ("valid", mask != 0), ("pad", mask == 0)):
        if not sel.any():
            continue
        g, r = got[sel], ref[sel]
        top2 = np.sort(r, -1)[:, -2:]
        gap = top2[:, 1] - top2[:, 0]
        agree = g.argmax(-1) == r.argmax(-1)
        out[lab] = {"n": int(sel.sum()), "max_abs": float(np.abs(g - r).max()),
                    "rel_max": float(np.abs(g - r).max() / np.abs(ref).max()),
                    "rms": float(np.sqrt(((g - r) ** 2).mean())),
                    "top1": float(agree.mean()),
                    "top1_gap>1e-2": float(agree[gap > 1e-2].mean()) if (gap > 1e-2).any() else 1.0,
                    "min_gap_of_miss": float(gap[~agree].max()) if (~agree).any() else None}
    print(json.dumps(out), flush=True)


def main():
    spec = bo.BERT_BASE
    w = bo.make_weights(spec, 0)
    r = BertReplica(w, max_batch=64, max_seq=512, n_workspaces=1)
    for B, S, min_len in [(1, 128, None), (4, 128, 16), (3, 200, 16)]:
        ids, mask = bo.make_inputs(spec, B, S, seed=77 + B + S, min_len=min_len)
        ref = bo.bert_forward(w, spec, ids, mask)
        emu = emulated(w, spec, ids, mask)
        out = torch.empty(B, S, spec.hidden, device="cuda")
        r.forward_async(torch.from_numpy(ids).cuda(), torch.from_numpy(mask).cuda(), out)
        torch.cuda.synchronize()
        stats(f"cuda  B{B} S{S}", out.cpu().numpy(), ref, mask)
        stats(f"emu   B{B} S{S}", emu, ref, mask)
    r.close()


if __name__ == "__main__":
    main()
