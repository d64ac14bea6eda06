"""Seeded synthetic weights and requests for benchmarks and warm-up (numpy PCG64).

The reference warms up and benchmarks on random inputs too
(/root/reference/src/utils/input_generator.hpp:22-56); BASELINE.json asks for random-init
weights of the named architecture.  These generators draw the same streams as the test
oracle's (checked by tests/test_synthetic.py) but are independent product code: the
benchmark never imports the oracle for its GPU arm."""
from __future__ import annotations

import numpy as np

from .bert import BertConfig, _tensor_order


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
