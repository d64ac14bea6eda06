"""fp32 CPU oracle of the arithmetic the reference codelet executes for BERT.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

What it restates
----------------
The reference codelet body is ``model->forward(ivalues)``
(/root/reference/src/core/starpu_setup.cpp:594-624, call at :610); the model is
the TorchScript trace of HF ``BertModel`` wrapped to return ``[0]`` =
``last_hidden_state`` (/root/reference/models/import_bert-base-uncased.py:8-25).
The arithmetic therefore lives in a third-party dependency that is NOT under
/root/reference: HuggingFace ``transformers`` (unpinned in the exporter,
``transformers>=4.37`` in /root/reference/client/requirements.txt:3; 5.5.0 in
this image) executed by LibTorch 2.2.2 (/root/reference/Dockerfile:50-51).
This file restates the published BertModel graph:

  embeddings  = LN(word[ids] + position[0..S) + token_type[0], eps=1e-12)
  per layer   : q,k,v = x W^T + b ; scores = q k^T / sqrt(d_h) + (1-m)*finfo.min
                p = softmax(scores) ; ctx = p v
                x = LN(ctx Wo^T + bo + x) ; h = gelu_erf(x W1^T + b1)
                x = LN(h W2^T + b2 + x)
  output      = x   (last_hidden_state; the pooler is never evaluated)

Pinning: ``oracle/make_golden.py`` imports HF ``BertModel`` in the build
container, loads the same seeded weights and commits input/output vectors under
``tests/golden/``; ``tests/test_oracle_golden.py`` checks this restatement
against them.  The reference repository itself holds no BERT golden vector
(SURVEY.md section 8c), so those HF-generated fixtures are the pin.

Weights use the HF state-dict key names so the same dict loads into
``BertModel`` (strict=False only for the unused pooler).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, asdict

import numpy as np
import torch


@dataclass(frozen=True)
class BertSpec:
    layers: int = 12
    hidden: int = 768
    heads: int = 12
    ffn: int = 3072
    vocab: int = 30522
    max_pos: int = 512
    type_vocab: int = 2
    ln_eps: float = 1e-12

    @property
    def head_dim(self) -> int:
        return self.hidden // self.heads

    def to_dict(self):
        return asdict(self)


BERT_BASE = BertSpec()
# A small spec with the same structure for fast CPU tests / committed fixtures.
BERT_TINY = BertSpec(layers=2, hidden=128, heads=2, ffn=512, vocab=1000, max_pos=512)


def weight_names(spec: BertSpec):
    """(name, shape, kind) in a fixed order; kind in {'w','b','g'}."""
    H, F = spec.hidden, spec.ffn
    out = [
        ("embeddings.word_embeddings.weight", (spec.vocab, H), "w"),
        ("embeddings.position_embeddings.weight", (spec.max_pos, H), "w"),
        ("embeddings.token_type_embeddings.weight", (spec.type_vocab, H), "w"),
        ("embeddings.LayerNorm.weight", (H,), "g"),
        ("embeddings.LayerNorm.bias", (H,), "b"),
    ]
    for l in range(spec.layers):
        p = f"encoder.layer.{l}."
        out += [
            (p + "attention.self.query.weight", (H, H), "w"),
            (p + "attention.self.query.bias", (H,), "b"),
            (p + "attention.self.key.weight", (H, H), "w"),
            (p + "attention.self.key.bias", (H,), "b"),
            (p + "attention.self.value.weight", (H, H), "w"),
            (p + "attention.self.value.bias", (H,), "b"),
            (p + "attention.output.dense.weight", (H, H), "w"),
            (p + "attention.output.dense.bias", (H,), "b"),
            (p + "attention.output.LayerNorm.weight", (H,), "g"),
            (p + "attention.output.LayerNorm.bias", (H,), "b"),
            (p + "intermediate.dense.weight", (F, H), "w"),
            (p + "intermediate.dense.bias", (F,), "b"),
            (p + "output.dense.weight", (H, F), "w"),
            (p + "output.dense.bias", (H,), "b"),
            (p + "output.LayerNorm.weight", (H,), "g"),
            (p + "output.LayerNorm.bias", (H,), "b"),
        ]
    return out


def make_weights(spec: BertSpec = BERT_BASE, seed: int = 0, w_std: float = 0.02):
    """Seeded random-init weights (numpy PCG64, stable across library versions).

    Distribution follows HF ``BertPreTrainedModel._init_weights`` (normal(0,
    initializer_range=0.02)) for matrices; biases and LayerNorm affine terms are
    drawn non-trivially (bias ~ N(0, 0.02), gamma ~ 1 + N(0, 0.02)) so that every
    bias/affine path of the kernels is exercised by parity tests.
    """
    rng = np.random.Generator(np.random.PCG64(seed))
    w = {}
    for name, shape, kind in weight_names(spec):
        a = rng.standard_normal(shape, dtype=np.float32)
        if kind == "w":
            a *= np.float32(w_std)
        elif kind == "b":
            a *= np.float32(0.02)
        else:
            a = np.float32(1.0) + a * np.float32(0.02)
        w[name] = a
    return w


def make_inputs(spec: BertSpec, batch: int, seq: int, seed: int = 1234,
                min_len: int | None = None):
    """Seeded ids ~ U[0, vocab) and a 1/0 attention mask with a padded tail.

    ``min_len=None`` -> all-ones mask.  Otherwise row r keeps L_r ~ U[min_len, seq]
    tokens (ids of the tail are 0, like the HF tokenizer's [PAD]).
    """
    rng = np.random.Generator(np.random.PCG64(seed))
    ids = rng.integers(0, spec.vocab, size=(batch, seq), dtype=np.int64)
    mask = np.ones((batch, seq), dtype=np.int64)
    if min_len is not None:
        lens = rng.integers(min_len, seq + 1, size=(batch,))
        for r, L in enumerate(lens):
            mask[r, L:] = 0
            ids[r, L:] = 0
    return ids, mask


def _ln(x: torch.Tensor, g: torch.Tensor, b: torch.Tensor, eps: float) -> torch.Tensor:
    mu = x.mean(-1, keepdim=True)
    var = ((x - mu) ** 2).mean(-1, keepdim=True)
    return (x - mu) / torch.sqrt(var + eps) * g + b


def _gelu_erf(x: torch.Tensor) -> torch.Tensor:
    return x * 0.5 * (1.0 + torch.erf(x / math.sqrt(2.0)))


def _t(w, name):
    a = w[name]
    return a if isinstance(a, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(a))


@torch.no_grad()
def bert_forward(w, spec: BertSpec, input_ids, attention_mask, *, return_layers: bool = False):
    """fp32 forward; returns last_hidden_state [B,S,H] as a numpy array.

    ``attention_mask``: any integer tensor; non-zero = keep (the reference warm-up
    feeds randint(0,10) masks, /root/reference/src/utils/input_generator.hpp:22-38).
    """
    ids = torch.as_tensor(np.asarray(input_ids)).long()
    keep = torch.as_tensor(np.asarray(attention_mask)) != 0
    B, S = ids.shape
    H, nh, dh = spec.hidden, spec.heads, spec.head_dim
    x = (_t(w, "embeddings.word_embeddings.weight")[ids]
         + _t(w, "embeddings.position_embeddings.weight")[:S][None]
         + _t(w, "embeddings.token_type_embeddings.weight")[0][None, None])
    x = _ln(x, _t(w, "embeddings.LayerNorm.weight"), _t(w, "embeddings.LayerNorm.bias"), spec.ln_eps)
    add_mask = torch.zeros(B, 1, 1, S, dtype=torch.float32)
    add_mask.masked_fill_(~keep[:, None, None, :], torch.finfo(torch.float32).min)
    layers = [x.numpy().copy()] if return_layers else None
    for l in range(spec.layers):
        p = f"encoder.layer.{l}."
        lin = lambda t, n: t @ _t(w, p + n + ".weight").T + _t(w, p + n + ".bias")
        q = lin(x, "attention.self.query").view(B, S, nh, dh).transpose(1, 2)
        k = lin(x, "attention.self.key").view(B, S, nh, dh).transpose(1, 2)
        v = lin(x, "attention.self.value").view(B, S, nh, dh).transpose(1, 2)
        s = q @ k.transpose(-1, -2) / math.sqrt(dh) + add_mask
        pr = torch.softmax(s, dim=-1)
        ctx = (pr @ v).transpose(1, 2).reshape(B, S, H)
        x = _ln(lin(ctx, "attention.output.dense") + x,
                _t(w, p + "attention.output.LayerNorm.weight"),
                _t(w, p + "attention.output.LayerNorm.bias"), spec.ln_eps)
        h = _gelu_erf(lin(x, "intermediate.dense"))
        x = _ln(lin(h, "output.dense") + x,
                _t(w, p + "output.LayerNorm.weight"),
                _t(w, p + "output.LayerNorm.bias"), spec.ln_eps)
        if return_layers:
            layers.append(x.numpy().copy())
    out = x.numpy()
    return (out, layers) if return_layers else out


class BertOracleModule(torch.nn.Module):
    """Same graph as ``bert_forward`` as an nn.Module, so it can be
    ``torch.jit.trace``d -- the stand-in for ``bert_libtorch.pt`` when timing the
    reference's CPU codelet path (TorchScript forward on the host cores);
    ``torch.jit.trace`` of HF BertModel itself fails under transformers 5.5.0
    (SURVEY.md section 7)."""

    def __init__(self, w, spec: BertSpec):
        super().__init__()
        self.spec = spec
        self.params = torch.nn.ParameterDict(
            {k.replace(".", "__"): torch.nn.Parameter(_t(w, k).clone(), requires_grad=False) for k in w})

    def _p(self, name):
        return self.params[name.replace(".", "__")]

    def forward(self, input_ids: torch.Tensor, attention_mask: torch.Tensor) -> torch.Tensor:
        spec = self.spec
        B, S = input_ids.shape[0], input_ids.shape[1]
        H, nh, dh = spec.hidden, spec.heads, spec.head_dim
        x = (self._p("embeddings.word_embeddings.weight")[input_ids]
             + self._p("embeddings.position_embeddings.weight")[:S].unsqueeze(0)
             + self._p("embeddings.token_type_embeddings.weight")[0])
        x = torch.nn.functional.layer_norm(
            x, (H,), self._p("embeddings.LayerNorm.weight"), self._p("embeddings.LayerNorm.bias"), spec.ln_eps)
        add_mask = (attention_mask == 0).to(torch.float32)[:, None, None, :] * torch.finfo(torch.float32).min
        for l in range(spec.layers):
            p = f"encoder.layer.{l}."
            lin = lambda t, n: torch.nn.functional.linear(t, self._p(p + n + ".weight"), self._p(p + n + ".bias"))
            q = lin(x, "attention.self.query").view(B, S, nh, dh).transpose(1, 2)
            k = lin(x, "attention.self.key").view(B, S, nh, dh).transpose(1, 2)
            v = lin(x, "attention.self.value").view(B, S, nh, dh).transpose(1, 2)
            s = q @ k.transpose(-1, -2) / math.sqrt(dh) + add_mask
            ctx = (torch.softmax(s, dim=-1) @ v).transpose(1, 2).reshape(B, S, H)
            x = torch.nn.functional.layer_norm(
                lin(ctx, "attention.output.dense") + x, (H,),
                self._p(p + "attention.output.LayerNorm.weight"),
                self._p(p + "attention.output.LayerNorm.bias"), spec.ln_eps)
            h = torch.nn.functional.gelu(lin(x, "intermediate.dense"))
            x = torch.nn.functional.layer_norm(
                lin(h, "output.dense") + x, (H,),
                self._p(p + "output.LayerNorm.weight"),
                self._p(p + "output.LayerNorm.bias"), spec.ln_eps)
        return x


def hf_bert_model(w, spec: BertSpec):
    """Build HF ``BertModel`` holding exactly ``w`` (build container only; used by
    make_golden.py to mint fixtures and by the CPU oracle-pin test when
    transformers is importable)."""
    from transformers import BertConfig, BertModel

    cfg = BertConfig(vocab_size=spec.vocab, hidden_size=spec.hidden, num_hidden_layers=spec.layers,
                     num_attention_heads=spec.heads, intermediate_size=spec.ffn,
                     max_position_embeddings=spec.max_pos, type_vocab_size=spec.type_vocab,
                     layer_norm_eps=spec.ln_eps, hidden_act="gelu",
                     hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
    m = BertModel(cfg, add_pooling_layer=False).eval()
    sd = {k: _t(w, k).clone() for k in w}
    missing, unexpected = m.load_state_dict(sd, strict=False)
    missing = [k for k in missing if "position_ids" not in k and "token_type_ids" not in k]
    assert not missing and not unexpected, (missing, unexpected)
    return m
