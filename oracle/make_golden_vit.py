"""Mint the ViT golden fixtures under tests/golden/ from the REAL third-party model
(torchvision VisionTransformer, CPU).  Build container only:

    python -m oracle.make_golden_vit

Inputs are regenerated from their seed (602 KB per image would bloat the repo); only the
seeds and the torchvision fp32 logits are stored.  Weights regenerate from the seed.

Fixtures
  vit_tiny.npz   VIT_TINY (64x64, 2 layers), B=3, input seed 21: logits [3,10]
  vit_l16.npz    ViT-L/16 224x224, B=2, input seed 22: logits [2,1000]
  MANIFEST_VIT.json   seeds, versions, sha256
"""
from __future__ import annotations

import hashlib
import json
import os

import numpy as np
import torch

from . import vit_oracle as vo

GOLD = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def main():
    import torchvision

    torch.set_num_threads(max(1, os.cpu_count() or 1))
    manifest = {"generator": "oracle/make_golden_vit.py", "torch": torch.__version__,
                "torchvision": torchvision.__version__, "numpy": np.__version__,
                "weights": "oracle.vit_oracle.make_weights(spec, seed=0)",
                "inputs": "oracle.vit_oracle.make_inputs(spec, batch, seed)", "files": {}}
    for name, spec, batch, seed in (("vit_tiny.npz", vo.VIT_TINY, 3, 21), ("vit_l16.npz", vo.VIT_L_16, 2, 22)):
        w = vo.make_weights(spec, 0)
        x = vo.make_inputs(spec, batch, seed)
        with torch.no_grad():
            logits = vo.torchvision_model(w, spec)(torch.from_numpy(x)).numpy()
        path = os.path.join(GOLD, name)
        np.savez_compressed(path, logits=logits, batch=batch, seed=seed, pixel_sum=np.float64(x.astype(np.float64).sum()))
        manifest["files"][name] = hashlib.sha256(open(path, "rb").read()).hexdigest()
        print(name, logits.shape, float(np.abs(logits).max()), np.sort(logits[0])[-3:])
    with open(os.path.join(GOLD, "MANIFEST_VIT.json"), "w") as f:
        json.dump(manifest, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
