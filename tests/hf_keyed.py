"""Test helper: an nn.Module TREE whose parameter names are exactly HuggingFace BertModel's
(`embeddings.word_embeddings.weight`, `encoder.layer.N.attention.self.query.weight`, ...), wrapped like the
reference's exporter wraps BertModel (attribute `bert`, /root/reference/models/import_bert-base-uncased.py:8-25),
so that `torch.jit.trace(...).save("bert_libtorch.pt")` yields an artifact with the key layout the served file has.
(HF BertModel itself does not trace under transformers 5.5; the graph here is the oracle's.)"""
import math

import torch
from torch import nn


class _SelfAttention(nn.Module):
    def __init__(self, H):
        super().__init__()
        self.query, self.key, self.value = nn.Linear(H, H), nn.Linear(H, H), nn.Linear(H, H)


class _SelfOutput(nn.Module):
    def __init__(self, i, o, eps):
        super().__init__()
        self.dense = nn.Linear(i, o)
        self.LayerNorm = nn.LayerNorm(o, eps=eps)


class _Attention(nn.Module):
    def __init__(self, H, eps):
        super().__init__()
        self.self = _SelfAttention(H)
        self.output = _SelfOutput(H, H, eps)


class _Intermediate(nn.Module):
    def __init__(self, H, F):
        super().__init__()
        self.dense = nn.Linear(H, F)


class _Layer(nn.Module):
    def __init__(self, spec):
        super().__init__()
        self.attention = _Attention(spec.hidden, spec.ln_eps)
        self.intermediate = _Intermediate(spec.hidden, spec.ffn)
        self.output = _SelfOutput(spec.ffn, spec.hidden, spec.ln_eps)


class _Encoder(nn.Module):
    def __init__(self, spec):
        super().__init__()
        self.layer = nn.ModuleList([_Layer(spec) for _ in range(spec.layers)])


class _Embeddings(nn.Module):
    def __init__(self, spec):
        super().__init__()
        self.word_embeddings = nn.Embedding(spec.vocab, spec.hidden)
        self.position_embeddings = nn.Embedding(spec.max_pos, spec.hidden)
        self.token_type_embeddings = nn.Embedding(spec.type_vocab, spec.hidden)
        self.LayerNorm = nn.LayerNorm(spec.hidden, eps=spec.ln_eps)
        self.register_buffer("position_ids", torch.arange(spec.max_pos).unsqueeze(0), persistent=True)


class HFKeyedBert(nn.Module):
    def __init__(self, spec):
        super().__init__()
        self.heads, self.head_dim = spec.heads, spec.hidden // spec.heads
        self.embeddings = _Embeddings(spec)
        self.encoder = _Encoder(spec)

    def forward(self, input_ids, attention_mask):
        B, S = input_ids.shape[0], input_ids.shape[1]
        e = self.embeddings
        x = e.word_embeddings(input_ids) + e.position_embeddings(e.position_ids[:, :S]) + e.token_type_embeddings.weight[0]
        x = e.LayerNorm(x)
        add_mask = (attention_mask == 0).to(torch.float32)[:, None, None, :] * torch.finfo(torch.float32).min
        for lyr in self.encoder.layer:
            a = lyr.attention
            split = lambda t: t.view(B, S, self.heads, self.head_dim).transpose(1, 2)  # noqa: E731
            q, k, v = split(a.self.query(x)), split(a.self.key(x)), split(a.self.value(x))
            s = q @ k.transpose(-1, -2) / math.sqrt(self.head_dim) + add_mask
            ctx = (torch.softmax(s, dim=-1) @ v).transpose(1, 2).reshape(B, S, -1)
            x = a.output.LayerNorm(a.output.dense(ctx) + x)
            h = torch.nn.functional.gelu(lyr.intermediate.dense(x))
            x = lyr.output.LayerNorm(lyr.output.dense(h) + x)
        return x


class BertWrapper(nn.Module):
    """The exporter's wrapper: state-dict keys carry the `bert.` prefix."""

    def __init__(self, spec, weights):
        super().__init__()
        self.bert = HFKeyedBert(spec)
        sd = {k: torch.from_numpy(v) for k, v in weights.items()}
        missing, unexpected = self.bert.load_state_dict(sd, strict=False)
        assert not unexpected and set(missing) <= {"embeddings.position_ids"}, (missing, unexpected)
        self.bert.eval()

    def forward(self, input_ids: torch.Tensor, attention_mask: torch.Tensor) -> torch.Tensor:
        return self.bert(input_ids, attention_mask)


def export_bert_torchscript(spec, weights, path):
    """What models/import_bert-base-uncased.py does: trace the wrapper on one example and save it."""
    m = BertWrapper(spec, weights).eval()
    ids = torch.randint(0, spec.vocab, (1, 8))
    with torch.no_grad():
        traced = torch.jit.trace(m, (ids, torch.ones_like(ids)), check_trace=False)
        traced.save(path)
    return m
