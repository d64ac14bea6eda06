"""CPU: the oracle restatement (oracle/bert_oracle.py, oracle/gather_scatter.c) against
the committed golden fixtures minted from HF BertModel (oracle/make_golden.py) and against
the known answers the reference's own tests hold for gather/scatter."""
import ctypes as C
import hashlib
import json
import os

import numpy as np
import pytest

from oracle import bert_oracle as bo

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _load(name):
    z = np.load(os.path.join(GOLD, name))
    return z, bo.BertSpec(**json.loads(str(z["spec"])))


def test_manifest_hashes():
    man = json.load(open(os.path.join(GOLD, "MANIFEST.json")))
    for name, sha in man["files"].items():
        assert hashlib.sha256(open(os.path.join(GOLD, name), "rb").read()).hexdigest() == sha, name


@pytest.mark.parametrize("name", ["bert_tiny_s32.npz", "bert_tiny_s128.npz"])
def test_oracle_matches_hf_tiny(name):
    z, spec = _load(name)
    w = bo.make_weights(spec, 0)
    out = bo.bert_forward(w, spec, z["ids"], z["mask"])
    keep = z["mask"] != 0
    # padded query rows are well defined too (HF computes them); compare everything
    assert np.allclose(out, z["out"], rtol=1e-4, atol=2e-5), np.abs(out - z["out"]).max()
    assert np.abs(out[keep] - z["out"][keep]).max() < 2e-5


@pytest.fixture(scope="module")
def base_weights():
    return bo.make_weights(bo.BERT_BASE, 0)


@pytest.mark.parametrize("name", ["bert_base_s128.npz", "bert_base_s128_full.npz"])
def test_oracle_matches_hf_base(name, base_weights):
    z, spec = _load(name)
    assert spec == bo.BERT_BASE
    out = bo.bert_forward(base_weights, spec, z["ids"], z["mask"])
    sub = out[:, ::8, :]
    assert np.allclose(sub, z["out_sub"], rtol=1e-4, atol=5e-5), np.abs(sub - z["out_sub"]).max()
    assert np.allclose(out.astype(np.float64).sum(-1), z["row_sums"], atol=2e-3)


def test_traceable_module_equals_function():
    import torch

    spec = bo.BERT_TINY
    w = bo.make_weights(spec, 0)
    ids, mask = bo.make_inputs(spec, 2, 24, seed=3, min_len=4)
    m = bo.BertOracleModule(w, spec).eval()
    with torch.no_grad():
        a = m(torch.from_numpy(ids), torch.from_numpy(mask)).numpy()
        ts = torch.jit.trace(m, (torch.from_numpy(ids), torch.from_numpy(mask)))
        b = ts(torch.from_numpy(ids), torch.from_numpy(mask)).numpy()
    ref = bo.bert_forward(w, spec, ids, mask)
    assert np.allclose(a, ref, rtol=1e-4, atol=2e-5)
    assert np.allclose(b, ref, rtol=1e-4, atol=2e-5)


def test_nonbinary_mask_means_keep():
    spec = bo.BERT_TINY
    w = bo.make_weights(spec, 0)
    ids, mask = bo.make_inputs(spec, 2, 16, seed=4, min_len=4)
    a = bo.bert_forward(w, spec, ids, mask)
    b = bo.bert_forward(w, spec, ids, mask * 9)
    assert np.array_equal(a, b)


# ---- gather / scatter known answers from the reference's tests -------------------
def test_gather_concat_known_answer(oracle_gs):
    """jobs {1,2}+{3,4} -> slot 1,2,3,4 and nx == 4
    (tests/unit/starpu/unit_starpu_task_runner_submission_and_slots.cpp:607-666)."""
    a, b = np.array([1, 2], np.float32), np.array([3, 4], np.float32)
    slot = np.zeros(4, np.float32)
    nx, by = C.c_size_t(), C.c_size_t()
    rc = oracle_gs.oracle_gather_concat((C.c_void_p * 2)(a.ctypes.data, b.ctypes.data),
                                        (C.c_size_t * 2)(8, 8), (C.c_size_t * 2)(2, 2), C.c_size_t(2),
                                        slot.ctypes.data_as(C.c_void_p), C.c_size_t(16), C.byref(nx),
                                        C.byref(by))
    assert rc == 0 and nx.value == 4 and by.value == 16
    assert slot.tolist() == [1, 2, 3, 4]
    # capacity overflow raises InputPoolCapacityException in the reference
    rc = oracle_gs.oracle_gather_concat((C.c_void_p * 2)(a.ctypes.data, b.ctypes.data),
                                        (C.c_size_t * 2)(8, 8), (C.c_size_t * 2)(2, 2), C.c_size_t(2),
                                        slot.ctypes.data_as(C.c_void_p), C.c_size_t(12), C.byref(nx),
                                        C.byref(by))
    assert rc == 1


def test_scatter_slice_known_answer(oracle_gs):
    """batch sizes 1+2 over 3 rows -> rows [0,1) and [1,3)
    (tests/unit/starpu/unit_starpu_task_runner_batching_and_internal.cpp:1514-1606)."""
    agg = np.arange(6, dtype=np.float32).reshape(3, 2)
    d0, d1 = np.zeros((1, 2), np.float32), np.zeros((2, 2), np.float32)
    got = C.c_int64()
    for off, bs, d in ((0, 1, d0), (1, 2, d1)):
        rc = oracle_gs.oracle_scatter_slice(agg.ctypes.data_as(C.c_void_p), C.c_size_t(3),
                                            C.c_size_t(agg.nbytes), C.c_int64(off), C.c_int64(bs),
                                            d.ctypes.data_as(C.c_void_p), C.byref(got))
        assert rc == 0 and got.value == bs
    assert d0.tolist() == [[0, 1]] and d1.tolist() == [[2, 3], [4, 5]]
    # slice past the end is clamped; offset beyond rows yields nothing
    d = np.zeros((2, 2), np.float32)
    oracle_gs.oracle_scatter_slice(agg.ctypes.data_as(C.c_void_p), C.c_size_t(3), C.c_size_t(agg.nbytes),
                                   C.c_int64(2), C.c_int64(2), d.ctypes.data_as(C.c_void_p), C.byref(got))
    assert got.value == 1 and d[0].tolist() == [4, 5]
    oracle_gs.oracle_scatter_slice(agg.ctypes.data_as(C.c_void_p), C.c_size_t(3), C.c_size_t(agg.nbytes),
                                   C.c_int64(3), C.c_int64(1), d.ctypes.data_as(C.c_void_p), C.byref(got))
    assert got.value == 0


def test_buffer_byte_size(oracle_gs):
    out = C.c_uint64()
    assert oracle_gs.oracle_buffer_byte_size(2, C.c_uint64(4096), C.c_uint64(8), C.byref(out)) == 0
    assert out.value == 32768
    assert oracle_gs.oracle_buffer_byte_size(5, C.c_uint64(77), C.c_uint64(4), C.byref(out)) == 0
    assert out.value == 4
    assert oracle_gs.oracle_buffer_byte_size(2, C.c_uint64(2**63), C.c_uint64(4), C.byref(out)) == 1
    assert oracle_gs.oracle_buffer_byte_size(9, C.c_uint64(1), C.c_uint64(4), C.byref(out)) == 2


def test_ragged_gather_scatter_roundtrip(oracle_gs):
    rng = np.random.Generator(np.random.PCG64(1))
    lens = np.array([5, 0, 16, 9], np.int32)
    seq = 16
    src = [rng.integers(1, 1000, size=max(L, 1), dtype=np.int64) for L in lens]
    ids = np.zeros((4, seq), np.int64)
    mask = np.zeros((4, seq), np.int64)
    PP = C.c_void_p * 4
    rc = oracle_gs.oracle_gather_pad_mask(PP(*[s.ctypes.data for s in src]), None,
                                          lens.ctypes.data_as(C.c_void_p), C.c_size_t(4), C.c_size_t(seq),
                                          C.c_int64(0), ids.ctypes.data_as(C.c_void_p),
                                          mask.ctypes.data_as(C.c_void_p))
    assert rc == 0
    assert (mask.sum(1) == lens).all()
    dst = [np.zeros(max(L, 1), np.int64) for L in lens]
    rc = oracle_gs.oracle_scatter_ragged(ids.ctypes.data_as(C.c_void_p), C.c_size_t(4), C.c_size_t(seq),
                                         C.c_size_t(8), lens.ctypes.data_as(C.c_void_p),
                                         PP(*[d.ctypes.data for d in dst]))
    assert rc == 0
    for s, d, L in zip(src, dst, lens):
        assert np.array_equal(s[:L], d[:L])


@pytest.mark.parametrize("name,spec_name", [("vit_tiny.npz", "VIT_TINY"), ("vit_l16.npz", "VIT_L_16")])
def test_vit_oracle_matches_torchvision(name, spec_name):
    """oracle/vit_oracle.py restatement vs the torchvision fp32 logits (oracle/make_golden_vit.py)."""
    import hashlib
    import json

    from oracle import vit_oracle as vo

    man = json.load(open(os.path.join(GOLD, "MANIFEST_VIT.json")))
    assert hashlib.sha256(open(os.path.join(GOLD, name), "rb").read()).hexdigest() == man["files"][name]
    spec = getattr(vo, spec_name)
    z = np.load(os.path.join(GOLD, name))
    if spec_name == "VIT_L_16" and (os.cpu_count() or 1) < 4:
        pytest.skip("ViT-L on the CPU needs a few cores")
    x = vo.make_inputs(spec, int(z["batch"]), int(z["seed"]))
    got = vo.vit_forward(vo.make_weights(spec, 0), spec, x)
    assert np.abs(got - z["logits"]).max() < 2e-5 * max(1.0, float(np.abs(z["logits"]).max()))
    assert (got.argmax(-1) == z["logits"].argmax(-1)).all()


def test_vit_im2col_is_conv_proj():
    """The patch embedding written as a GEMM over im2col rows equals nn.Conv2d(stride=patch)."""
    import torch

    from oracle import vit_oracle as vo

    spec = vo.VIT_TINY
    w = vo.make_weights(spec, 3)
    x = torch.from_numpy(vo.make_inputs(spec, 2, seed=9))
    conv = torch.nn.functional.conv2d(x, torch.from_numpy(w["conv_proj.weight"]), torch.from_numpy(w["conv_proj.bias"]),
                                      stride=spec.patch)
    ref = conv.reshape(2, spec.hidden, -1).permute(0, 2, 1).reshape(-1, spec.hidden)
    got = vo.im2col(spec, x) @ torch.from_numpy(w["conv_proj.weight"]).reshape(spec.hidden, -1).T + \
        torch.from_numpy(w["conv_proj.bias"])
    assert torch.allclose(got, ref, atol=1e-5)
