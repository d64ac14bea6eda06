"""CPU: the reference's CPU codelet body (oracle/_ref/libref_cpu_codelet.so = the reference's own tensor_builder.cpp
+ LibTorch, see oracle/ref_cpu_codelet.cpp) against the Python oracle and the committed HF golden vectors: this pins
oracle/bert_oracle.py to what the reference path itself computes here, not only to fixtures minted elsewhere."""
import os

import numpy as np
import pytest

from oracle import bert_oracle as bo
from oracle import ref_cpu_codelet as rc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.skipif(not rc.available(), reason="oracle/_ref not built (needs /root/reference: make -C oracle _ref)")


def test_reference_tensor_builder_forward_matches_oracle(tmp_path):
    spec = bo.BERT_TINY
    w = bo.make_weights(spec, 0)
    path = rc.export_bert_torchscript(w, spec, str(tmp_path / "bert_libtorch.pt"), batch=2, seq=32)
    ids, mask = bo.make_inputs(spec, 3, 32, seed=5, min_len=7)  # another batch size than the traced example
    out = np.full((3, 32, spec.hidden), np.nan, dtype=np.float32)
    with rc.RefCpuCodelet(path, threads=2) as cl:
        ms = cl.run([ids, mask], out, warmup=1, iters=2)
    assert len(ms) == 2 and all(t > 0 for t in ms)
    ref = bo.bert_forward(w, spec, ids, mask)
    assert np.isfinite(out).all() and np.abs(out - ref).max() < 1e-5


def test_reference_codelet_reproduces_the_hf_golden(tmp_path):
    g = np.load(os.path.join(ROOT, "tests", "golden", "bert_tiny_s32.npz"))
    spec = bo.BERT_TINY
    w = bo.make_weights(spec, 0)  # tests/golden/MANIFEST.json: weights = make_weights(spec, seed=0)
    ids, mask = g["ids"], g["mask"]
    path = rc.export_bert_torchscript(w, spec, str(tmp_path / "m.pt"), batch=ids.shape[0], seq=ids.shape[1])
    out = np.empty(g["out"].shape, dtype=np.float32)
    with rc.RefCpuCodelet(path, threads=2) as cl:
        cl.run([ids, mask], out)
    assert np.abs(out - g["out"]).max() < 2e-5


def test_reference_copy_output_checks_the_buffer_size(tmp_path):
    """copy_output_to_buffer (tensor_builder.cpp:161-189) refuses a buffer whose byte size differs."""
    spec = bo.BERT_TINY
    w = bo.make_weights(spec, 0)
    path = rc.export_bert_torchscript(w, spec, str(tmp_path / "m.pt"), batch=1, seq=16)
    ids, mask = bo.make_inputs(spec, 1, 16, seed=2)
    out = np.empty((1, 16, spec.hidden + 1), dtype=np.float32)
    with rc.RefCpuCodelet(path) as cl, pytest.raises(RuntimeError, match="Output buffer size mismatch"):
        cl.run([ids, mask], out)
