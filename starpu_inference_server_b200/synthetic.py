"""Seeded synthetic weights and requests for benchmarks and warm-up (numpy PCG64).

The reference warms up and benchmarks on random inputs too
(/root/reference/src/utils/input_generator.hpp:22-56); BASELINE.json asks for random-init
weights of the named architecture.  These generators draw the same streams as the test
oracle's (checked by tests/test_synthetic.py) but are independent product code: the
benchmark never imports the oracle for its GPU arm."""
from __future__ import annotations

import numpy as np

import math

from .bert import BertConfig, _tensor_order
from .vit import VitConfig
from .vit import _tensor_order as _vit_tensor_order


def _shape_kind(name: str, cfg: BertConfig):
    H, F = cfg.hidden, cfg.ffn
    if name.endswith("word_embeddings.weight"):
        return (cfg.vocab, H), "w"
    if name.endswith("position_embeddings.weight"):
        return (cfg.max_pos, H), "w"
    if name.endswith("token_type_embeddings.weight"):
        return (cfg.type_vocab, H), "w"
    if "LayerNorm" in name:
        return (H,), "g" if name.endswith("weight") else "b"
    if "intermediate.dense" in name:
        return ((F, H), "w") if name.endswith("weight") else ((F,), "b")
    if ".output.dense" in name and "attention" not in name:
        return ((H, F), "w") if name.endswith("weight") else ((H,), "b")
    return ((H, H), "w") if name.endswith("weight") else ((H,), "b")


def make_weights(cfg: BertConfig = BertConfig(), seed: int = 0, w_std: float = 0.02):
    """HF-keyed fp32 state: matrices N(0, 0.02), biases N(0, 0.02), LN gamma 1+N(0, 0.02)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    out = {}
    for name in _tensor_order(cfg.layers):
        shape, kind = _shape_kind(name, cfg)
        a = rng.standard_normal(shape, dtype=np.float32)
        if kind == "w":
            a *= np.float32(w_std)
        elif kind == "b":
            a *= np.float32(0.02)
        else:
            a = np.float32(1.0) + a * np.float32(0.02)
        out[name] = a
    return out


def make_requests(cfg: BertConfig, batch: int, seq: int, seed: int = 1234, min_len: int | None = None):
    """ids ~ U[0, vocab) int64 [batch, seq] and a 1/0 mask (padded tail when min_len is set)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    ids = rng.integers(0, cfg.vocab, size=(batch, seq), dtype=np.int64)
    mask = np.ones((batch, seq), dtype=np.int64)
    if min_len is not None:
        lens = rng.integers(min_len, seq + 1, size=(batch,))
        for r, L in enumerate(lens):
            mask[r, L:] = 0
            ids[r, L:] = 0
    return ids, mask


def make_vit_weights(cfg: VitConfig = VitConfig(), seed: int = 0):
    """torchvision-keyed fp32 ViT state: projection / MLP matrices N(0, (0.5/sqrt(fan_in))^2) so activations keep
    O(1) scale through the pre-LN layers, head N(0, 0.05), biases N(0, 0.02), LayerNorm gamma 1+N(0, 0.02)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    H, M, S = cfg.hidden, cfg.mlp, cfg.seq
    shapes = {"conv_proj.weight": (H, cfg.channels, cfg.patch, cfg.patch), "conv_proj.bias": (H,),
              "class_token": (1, 1, H), "encoder.pos_embedding": (1, S, H), "encoder.ln.weight": (H,),
              "encoder.ln.bias": (H,), "heads.head.weight": (cfg.classes, H), "heads.head.bias": (cfg.classes,)}
    out = {}
    for name in _vit_tensor_order(cfg.layers):
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


def make_images(cfg: VitConfig, batch: int, seed: int = 1234):
    """pixels ~ U[0,1) fp32 [batch, C, H, W] (the reference's input generator draws rand(), input_generator.hpp:52-56)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    return rng.random((batch, cfg.channels, cfg.image, cfg.image), dtype=np.float32)
