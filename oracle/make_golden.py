"""Mint the golden fixtures under tests/golden/ from the REAL third-party model.

Run in the build container only (needs HuggingFace transformers, CPU):

    python -m oracle.make_golden

The arithmetic of the reference codelet is HF ``BertModel`` executed by LibTorch
(see oracle/bert_oracle.py header).  This script loads the seeded weights of
``oracle.bert_oracle.make_weights`` into HF ``BertModel`` and records its fp32
CPU outputs; the oracle restatement and the CUDA path are both checked against
these files.  Weights are NOT stored (438 MB): they regenerate from the seed.

Fixtures
  bert_tiny_s32.npz      BERT_TINY, B=3 S=32 ragged mask: ids, mask, full output
  bert_tiny_s128.npz     BERT_TINY, B=2 S=128 all-ones mask: ids, mask, full output
  bert_base_s128.npz     BERT_BASE, B=2 S=128 ragged mask: ids, mask, output at token
                         positions 0,8,..,120 (+ per-token fp64 sums of every row)
  bert_base_s128_full.npz BERT_BASE, B=1 S=128 all-ones: same subsampling
  MANIFEST.json          seeds, specs, library versions, sha256 of every file
"""
from __future__ import annotations

import hashlib
import json
import os

import numpy as np
import torch

from . import bert_oracle as bo

GOLD = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _hf_out(spec, w, ids, mask):
    m = bo.hf_bert_model(w, spec)
    with torch.no_grad():
        return m(input_ids=torch.from_numpy(ids), attention_mask=torch.from_numpy(mask),
                 return_dict=False)[0].numpy()


def main():
    import transformers

    os.makedirs(GOLD, exist_ok=True)
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    manifest = {"generator": "oracle/make_golden.py", "torch": torch.__version__,
                "transformers": transformers.__version__, "numpy": np.__version__,
                "weights": "oracle.bert_oracle.make_weights(spec, seed=0)", "files": {}}

    def save(name, **arrs):
        path = os.path.join(GOLD, name)
        np.savez_compressed(path, **arrs)
        manifest["files"][name] = hashlib.sha256(open(path, "rb").read()).hexdigest()

    w_tiny = bo.make_weights(bo.BERT_TINY, 0)
    ids, mask = bo.make_inputs(bo.BERT_TINY, 3, 32, seed=11, min_len=5)
    save("bert_tiny_s32.npz", ids=ids, mask=mask, out=_hf_out(bo.BERT_TINY, w_tiny, ids, mask),
         spec=json.dumps(bo.BERT_TINY.to_dict()))
    ids, mask = bo.make_inputs(bo.BERT_TINY, 2, 128, seed=12)
    save("bert_tiny_s128.npz", ids=ids, mask=mask, out=_hf_out(bo.BERT_TINY, w_tiny, ids, mask),
         spec=json.dumps(bo.BERT_TINY.to_dict()))

    w = bo.make_weights(bo.BERT_BASE, 0)
    for name, (b, seed, min_len) in {"bert_base_s128.npz": (2, 1234, 16),
                                     "bert_base_s128_full.npz": (1, 1235, None)}.items():
        ids, mask = bo.make_inputs(bo.BERT_BASE, b, 128, seed=seed, min_len=min_len)
        out = _hf_out(bo.BERT_BASE, w, ids, mask)
        save(name, ids=ids, mask=mask, pos=np.arange(0, 128, 8), out_sub=out[:, ::8, :].copy(),
             row_sums=out.astype(np.float64).sum(-1), spec=json.dumps(bo.BERT_BASE.to_dict()))

    with open(os.path.join(GOLD, "MANIFEST.json"), "w") as f:
        json.dump(manifest, f, indent=1, sort_keys=True)
    print("wrote", sorted(manifest["files"]))


if __name__ == "__main__":
    main()
