"""fp32 CPU oracle of the arithmetic the reference codelet executes for ViT.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

The reference serves ``torch.jit.script(torchvision.models.vit_l_16(...))``
(/root/reference/models/import_vit.py:10-13,51-55; config models/vit_l_16.yml:13-16:
input fp32 [1,3,224,224] -> output fp32 [1,1000]) through ``model->forward``
(/root/reference/src/core/starpu_setup.cpp:610).  The arithmetic lives in torchvision
(pinned 0.17.2 in the reference CI, .github/workflows/ci.yml:642; 0.26.0 in this image),
which is not under /root/reference; this file restates the published
``VisionTransformer`` graph:

  patches  = conv_proj(x) as a GEMM over [B*196, C*P*P] x [H, C*P*P]^T + bias
  tokens   = cat(class_token, patches) + pos_embedding
  per layer (pre-LN): y = LN1(x); qkv = y Win^T + bin; ctx = softmax(q k^T / sqrt(d)) v
                      x = x + ctx Wo^T + bo; z = LN2(x); x = x + W2 gelu_erf(W1 z + b1) + b2
  logits   = head(LN(x)[:, 0])           (LayerNorm eps 1e-6)

Pinning: ``oracle/make_golden_vit.py`` loads the same seeded weights into torchvision's
``VisionTransformer`` and commits input/output vectors under ``tests/golden/``;
``tests/test_oracle_golden.py`` checks this restatement against them.  The reference holds
no ViT golden vector, so those fixtures are the pin.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np
import torch


@dataclass(frozen=True)
class VitSpec:
    layers: int = 24
    hidden: int = 1024
    heads: int = 16
    mlp: int = 4096
    patch: int = 16
    image: int = 224
    channels: int = 3
    classes: int = 1000
    ln_eps: float = 1e-6

    @property
    def n_patches(self) -> int:
        return (self.image // self.patch) ** 2

    @property
    def seq(self) -> int:
        return self.n_patches + 1

    @property
    def patch_dim(self) -> int:
        return self.channels * self.patch * self.patch


VIT_L_16 = VitSpec()
VIT_B_16 = VitSpec(layers=12, hidden=768, heads=12, mlp=3072)
# small spec with the same structure (head_dim 64) for fast CPU tests / committed fixtures
VIT_TINY = VitSpec(layers=2, hidden=128, heads=2, mlp=256, patch=16, image=64, classes=10)


def weight_names(spec: VitSpec):
    """(name, shape, kind) in the order spu_vit_replica_create expects."""
    H, M = spec.hidden, spec.mlp
    out = [
        ("conv_proj.weight", (H, spec.channels, spec.patch, spec.patch), "w"),
        ("conv_proj.bias", (H,), "b"),
        ("class_token", (1, 1, H), "w"),
        ("encoder.pos_embedding", (1, spec.seq, H), "w"),
    ]
    for l in range(spec.layers):
        p = f"encoder.layers.encoder_layer_{l}."
        out += [
            (p + "ln_1.weight", (H,), "g"), (p + "ln_1.bias", (H,), "b"),
            (p + "self_attention.in_proj_weight", (3 * H, H), "w"),
            (p + "self_attention.in_proj_bias", (3 * H,), "b"),
            (p + "self_attention.out_proj.weight", (H, H), "w"),
            (p + "self_attention.out_proj.bias", (H,), "b"),
            (p + "ln_2.weight", (H,), "g"), (p + "ln_2.bias", (H,), "b"),
            (p + "mlp.0.weight", (M, H), "w"), (p + "mlp.0.bias", (M,), "b"),
            (p + "mlp.3.weight", (H, M), "w"), (p + "mlp.3.bias", (H,), "b"),
        ]
    out += [("encoder.ln.weight", (H,), "g"), ("encoder.ln.bias", (H,), "b"),
            ("heads.head.weight", (spec.classes, H), "w"), ("heads.head.bias", (spec.classes,), "b")]
    return out


def make_weights(spec: VitSpec = VIT_L_16, seed: int = 0, w_std: float = 0.02):
    """Seeded random-init weights (numpy PCG64).  Matrices ~ N(0, 0.02) except the MLP /
    projection matrices, drawn with std 1/sqrt(fan_in)/2 so that activations keep O(1) scale
    through 24 pre-LN layers (torchvision's own init is xavier); biases ~ N(0, 0.02),
    LayerNorm gamma ~ 1 + N(0, 0.02), head ~ N(0, 0.05) so logits are well separated."""
    rng = np.random.Generator(np.random.PCG64(seed))
    w = {}
    for name, shape, kind in weight_names(spec):
        a = rng.standard_normal(shape, dtype=np.float32)
        if kind == "w":
            if name.startswith("heads."):
                a *= np.float32(0.05)
            elif a.ndim == 2 or name == "conv_proj.weight":
                fan_in = int(np.prod(shape[1:]))
                a *= np.float32(0.5 / math.sqrt(fan_in))
            else:
                a *= np.float32(w_std)
        elif kind == "b":
            a *= np.float32(0.02)
        else:
            a = np.float32(1.0) + a * np.float32(0.02)
        w[name] = a
    return w


def make_inputs(spec: VitSpec, batch: int, seed: int = 1234):
    """Seeded pixels ~ U[0,1) fp32 [B,C,H,W] (input_generator.hpp:52-56 draws rand())."""
    rng = np.random.Generator(np.random.PCG64(seed))
    return rng.random((batch, spec.channels, spec.image, spec.image), dtype=np.float32)


def im2col(spec: VitSpec, x: torch.Tensor) -> torch.Tensor:
    """[B,C,H,W] -> [B*n_patches, C*P*P]; row (b, py*nw+px), column c*P*P + i*P + j."""
    B, P, n = x.shape[0], spec.patch, spec.image // spec.patch
    x = x.reshape(B, spec.channels, n, P, n, P).permute(0, 2, 4, 1, 3, 5)
    return x.reshape(B * n * n, spec.patch_dim)


def vit_forward(w, spec: VitSpec, pixels: np.ndarray) -> np.ndarray:
    """fp32 restatement; returns logits [B, classes]."""
    t = {k: torch.from_numpy(np.ascontiguousarray(v)) for k, v in w.items()}
    x = torch.from_numpy(np.ascontiguousarray(pixels, dtype=np.float32))
    B, H, nh, S = x.shape[0], spec.hidden, spec.heads, spec.seq
    d = H // nh
    ln = lambda v, p: torch.nn.functional.layer_norm(v, (H,), t[p + ".weight"], t[p + ".bias"], spec.ln_eps)
    with torch.no_grad():
        patches = im2col(spec, x) @ t["conv_proj.weight"].reshape(H, -1).T + t["conv_proj.bias"]
        x = torch.cat([t["class_token"].expand(B, 1, H), patches.view(B, spec.n_patches, H)], 1)
        x = x + t["encoder.pos_embedding"]
        for l in range(spec.layers):
            p = f"encoder.layers.encoder_layer_{l}."
            y = ln(x, p + "ln_1")
            qkv = y @ t[p + "self_attention.in_proj_weight"].T + t[p + "self_attention.in_proj_bias"]
            q, k, v = (a.view(B, S, nh, d).transpose(1, 2) for a in qkv.split(H, dim=-1))
            pr = torch.softmax(q @ k.transpose(-1, -2) / math.sqrt(d), -1)
            ctx = (pr @ v).transpose(1, 2).reshape(B, S, H)
            x = x + ctx @ t[p + "self_attention.out_proj.weight"].T + t[p + "self_attention.out_proj.bias"]
            z = ln(x, p + "ln_2")
            h = torch.nn.functional.gelu(z @ t[p + "mlp.0.weight"].T + t[p + "mlp.0.bias"])
            x = x + h @ t[p + "mlp.3.weight"].T + t[p + "mlp.3.bias"]
        x = ln(x, "encoder.ln")[:, 0]
        return (x @ t["heads.head.weight"].T + t["heads.head.bias"]).numpy()


def torchvision_model(w, spec: VitSpec):
    """The third-party graph itself with the oracle's weights (golden generation only)."""
    from torchvision.models.vision_transformer import VisionTransformer

    m = VisionTransformer(image_size=spec.image, patch_size=spec.patch, num_layers=spec.layers,
                          num_heads=spec.heads, hidden_dim=spec.hidden, mlp_dim=spec.mlp,
                          num_classes=spec.classes).eval()
    missing, unexpected = m.load_state_dict({k: torch.from_numpy(v) for k, v in w.items()}, strict=True)
    assert not missing and not unexpected
    return m
