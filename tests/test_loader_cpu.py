"""CPU: the C++ TorchScript weight loader (libspu_loader.so, include/spu_loader.h) against artifacts exported here the
way the reference's scripts export them (models/import_bert-base-uncased.py: trace of a wrapper whose attribute is
`bert`; models/import_vit.py: torch.jit.script of torchvision's VisionTransformer).  Every tensor the loader hands to
spu_*_replica_create must equal the exported weights bit for bit, in the documented order, and the inferred descriptor
must equal the architecture.  (Creating replicas needs a GPU: tests/test_gpu_loader.py.)"""
import ctypes as C
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def ld(built_libs):
    from starpu_inference_server_b200 import loader

    return loader


def test_loader_library_exports_every_declared_symbol(ld):
    import re

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "spu_loader.h")).read()
    declared = set(re.findall(r"\b(spu_loader_\w+)\s*\(", hdr))
    assert declared == set(ld.SYMBOLS)
    L = C.CDLL(ld.LIB_PATH)
    for name in declared:
        assert getattr(L, name) is not None


def test_hot_path_libraries_do_not_link_torch(built_libs):
    """LibTorch stays out of the forward path: only libspu_loader.so may depend on it."""
    import subprocess

    for lib in built_libs:
        needed = subprocess.run(["readelf", "-d", lib], capture_output=True, text=True).stdout
        links_torch = "libtorch" in needed or "libc10" in needed
        assert links_torch == lib.endswith("libspu_loader.so"), lib


def test_bert_torchscript_state_roundtrip(ld, tmp_path):
    from oracle import bert_oracle as bo
    from starpu_inference_server_b200.bert import _tensor_order

    import hf_keyed

    spec = bo.BERT_TINY
    w = bo.make_weights(spec, 3)
    path = str(tmp_path / "bert_libtorch.pt")
    wrapper = hf_keyed.export_bert_torchscript(spec, w, path)
    # the exported graph is the oracle's graph (so the artifact is a faithful stand-in for the served file)
    ids, mask = bo.make_inputs(spec, 2, 24, seed=1, min_len=5)
    with torch.no_grad():
        got = torch.jit.load(path)(torch.from_numpy(ids), torch.from_numpy(mask)).numpy()
    assert np.abs(got - bo.bert_forward(w, spec, ids, mask)).max() < 1e-5
    assert all(k.startswith("bert.") for k in wrapper.state_dict())
    with ld.TorchScriptModule(path, ld.KIND_BERT) as m:
        d = m.desc()
        assert (d.layers, d.hidden, d.heads, d.ffn, d.vocab, d.max_pos, d.type_vocab) == (
            spec.layers, spec.hidden, spec.heads, spec.ffn, spec.vocab, spec.max_pos, spec.type_vocab)
        assert abs(d.ln_eps - 1e-12) < 1e-18
        ts = m.tensors()
        assert [n for n, _ in ts] == _tensor_order(spec.layers)
        for name, a in ts:
            assert np.array_equal(a, w[name].reshape(-1)), name


def test_vit_torchscript_state_roundtrip(ld, tmp_path):
    from oracle import vit_oracle as vo
    from starpu_inference_server_b200.vit import _tensor_order

    spec = vo.VIT_TINY
    w = vo.make_weights(spec, 2)
    model = vo.torchvision_model(w, spec)
    path = str(tmp_path / "vit_tiny.pt")
    torch.jit.script(model).save(path)  # models/import_vit.py:51-55, mode "script"
    with ld.TorchScriptModule(path, ld.KIND_VIT) as m:
        d = m.desc()
        assert (d.layers, d.hidden, d.heads, d.mlp, d.patch, d.image, d.channels, d.classes) == (
            spec.layers, spec.hidden, spec.heads, spec.mlp, spec.patch, spec.image, 3, spec.classes)
        ts = m.tensors()
        assert [n for n, _ in ts] == _tensor_order(spec.layers)
        sd = model.state_dict()
        for name, a in ts:
            assert np.array_equal(a, sd[name].numpy().reshape(-1)), name


def test_loader_errors(ld, tmp_path):
    with pytest.raises(ld.LoaderError, match="loader_open"):
        ld.TorchScriptModule(str(tmp_path / "missing.pt"), ld.KIND_BERT)

    class Toy(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.lin = torch.nn.Linear(4, 4)

        def forward(self, x):
            return self.lin(x)

    p = str(tmp_path / "toy.pt")
    torch.jit.script(Toy()).save(p)
    with pytest.raises(ld.LoaderError, match="not a BertModel"):
        ld.TorchScriptModule(p, ld.KIND_BERT)
    with pytest.raises(ld.LoaderError, match="not a torchvision ViT"):
        ld.TorchScriptModule(p, ld.KIND_VIT)


def test_heads_and_eps_come_from_the_artifact_not_from_a_guess(ld, tmp_path):
    """hidden / 64 heads and a per-family eps constant are right for BERT-base / ViT-B/L and silently
    wrong otherwise.  The loader reads both from the TorchScript artifact (graph constants, module attributes), a
    caller may override them (models/*.yml), and a model the attention kernels cannot serve (head_dim != 64) is
    refused at load time -- the reference's torch::jit::load serves such a module, this path must not pretend to."""
    import dataclasses

    import torchvision

    import hf_keyed
    from oracle import bert_oracle as bo

    # traced BERT with a non-default epsilon: read from the aten::layer_norm constants of the forward graph
    spec = dataclasses.replace(bo.BERT_TINY, ln_eps=1e-5)
    path = str(tmp_path / "bert_eps.pt")
    hf_keyed.export_bert_torchscript(spec, bo.make_weights(spec, 4), path)
    with ld.TorchScriptModule(path, ld.KIND_BERT) as m:
        d = m.desc()
        assert d.heads == spec.heads and d.ln_eps == pytest.approx(1e-5, rel=1e-6)
    with ld.TorchScriptModule(path, ld.KIND_BERT, ln_eps=3e-6) as m:  # explicit value wins
        assert m.desc().ln_eps == pytest.approx(3e-6, rel=1e-6)
    # same hidden size split into 4 heads of 32: hidden / 64 would have said 2 and computed the wrong attention
    spec32 = dataclasses.replace(bo.BERT_TINY, heads=4)
    path32 = str(tmp_path / "bert_h32.pt")
    hf_keyed.export_bert_torchscript(spec32, bo.make_weights(spec32, 4), path32)
    with pytest.raises(ld.LoaderError, match="4 heads of 32"):
        ld.TorchScriptModule(path32, ld.KIND_BERT)
    # scripted torchvision ViT with 2 heads of 80 (the ViT-H/14 head size): `num_heads` attribute, refused
    vit = torchvision.models.vision_transformer.VisionTransformer(
        image_size=32, patch_size=16, num_layers=1, num_heads=2, hidden_dim=160, mlp_dim=320, num_classes=10).eval()
    pv = str(tmp_path / "vit_h80.pt")
    torch.jit.script(vit).save(pv)
    with pytest.raises(ld.LoaderError, match="2 heads of 80.*head_dim 64"):
        ld.TorchScriptModule(pv, ld.KIND_VIT)
    # the family default is read, not assumed
    vit64 = torchvision.models.vision_transformer.VisionTransformer(
        image_size=32, patch_size=16, num_layers=1, num_heads=2, hidden_dim=128, mlp_dim=256, num_classes=10).eval()
    p64 = str(tmp_path / "vit_h64.pt")
    torch.jit.script(vit64).save(p64)
    with ld.TorchScriptModule(p64, ld.KIND_VIT) as m:
        d = m.desc()
        assert d.heads == 2 and d.ln_eps == pytest.approx(1e-6, rel=1e-6)
