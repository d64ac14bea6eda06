"""GPU: ViT forward through the C-ABI against the torchvision golden logits and the fp32
oracle (tolerance: BASELINE.json bf16 gate, 2e-2 max relative error, 100 % top-1)."""
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
BF16_TOL = 2e-2


@pytest.fixture(scope="module")
def spu(built_libs):
    from starpu_inference_server_b200 import capi

    assert torch.cuda.is_available() and capi.lib().spu_device_is_sm100(0) == 1
    return capi


def _run(rep, x, classes, ws=0):
    out = torch.full((x.shape[0], classes), float("nan"), device="cuda")
    rep.forward_async(torch.from_numpy(x).cuda(), out, ws=ws)
    torch.cuda.synchronize()
    return out.cpu().numpy()


def _check(got, ref):
    assert np.isfinite(got).all()
    err = float(np.abs(got - ref).max() / np.abs(ref).max())
    assert err < BF16_TOL, err
    assert (got.argmax(-1) == ref.argmax(-1)).all()
    return err


@pytest.fixture(scope="module")
def tiny(spu):
    from oracle import vit_oracle as vo
    from starpu_inference_server_b200.vit import VitReplica

    w = vo.make_weights(vo.VIT_TINY, 0)
    r = VitReplica(w, max_batch=9, n_workspaces=2)
    yield vo, w, r
    r.close()


def test_config_inferred_from_state(tiny):
    vo, w, r = tiny
    c = r.cfg
    assert (c.layers, c.hidden, c.heads, c.mlp, c.patch, c.image, c.classes) == (2, 128, 2, 256, 16, 64, 10)


def test_tiny_vs_torchvision_golden(tiny):
    vo, w, r = tiny
    z = np.load(os.path.join(GOLD, "vit_tiny.npz"))
    x = vo.make_inputs(vo.VIT_TINY, int(z["batch"]), int(z["seed"]))
    assert abs(float(x.astype(np.float64).sum()) - float(z["pixel_sum"])) < 1e-6
    _check(_run(r, x, 10), z["logits"])


@pytest.mark.parametrize("batch", [1, 4, 9])
def test_tiny_vs_oracle_and_workspaces(tiny, batch):
    vo, w, r = tiny
    x = vo.make_inputs(vo.VIT_TINY, batch, seed=100 + batch)
    ref = vo.vit_forward(w, vo.VIT_TINY, x)
    a = _run(r, x, 10, ws=0)
    _check(a, ref)
    assert (a == _run(r, x, 10, ws=1)).all()      # arenas are interchangeable, bit for bit
    if batch > 1:                                  # images are independent of their batch mates
        assert (a[1:2] == _run(r, x[1:2].copy(), 10)).all()


def test_capacity_and_argument_errors(tiny, spu):
    vo, w, r = tiny
    x = torch.zeros(10, 3, 64, 64, device="cuda")
    out = torch.zeros(10, 10, device="cuda")
    with pytest.raises(spu.SpuError) as e:
        r.forward_async(x, out)
    assert e.value.status == spu.SPU_ERR_CAPACITY
    with pytest.raises(spu.SpuError):
        r.forward_async(x[:1], out[:1], ws=5)
    assert spu.lib().spu_vit_input_bytes(r._h, 3) == 3 * 3 * 64 * 64 * 4
    assert spu.lib().spu_vit_output_bytes(r._h, 3) == 3 * 10 * 4
    assert spu.lib().spu_bert_output_bytes(r._h, 3, 8) == 0  # not a BERT replica


def test_vit_l16_vs_torchvision_golden(spu):
    from oracle import vit_oracle as vo
    from starpu_inference_server_b200.vit import VitReplica

    z = np.load(os.path.join(GOLD, "vit_l16.npz"))
    w = vo.make_weights(vo.VIT_L_16, 0)
    r = VitReplica(w, max_batch=8)
    try:
        x = vo.make_inputs(vo.VIT_L_16, int(z["batch"]), int(z["seed"]))
        err = _check(_run(r, x, 1000), z["logits"])
        print(f"ViT-L/16 B=2: rel err {err:.3e}")
        # a different batch composition must give the same logits for the same image
        x5 = np.concatenate([vo.make_inputs(vo.VIT_L_16, 3, seed=5), x], 0)
        got5 = _run(r, x5, 1000)
        assert np.abs(got5[3:] - z["logits"]).max() / np.abs(z["logits"]).max() < BF16_TOL
    finally:
        r.close()


def test_vit_through_the_task_path(tiny):
    """[1,3,64,64] requests -> host gather -> task -> codelet -> C-ABI -> D2H -> slices."""
    from starpu_inference_server_b200 import hostapi

    vo, w, r = tiny
    x = vo.make_inputs(vo.VIT_TINY, 5, seed=77)
    direct = _run(r, x, 10, ws=1)
    with hostapi.Server(inputs=[("input", "f32", [1, 3, 64, 64])], outputs=[("output", "f32", [1, 10])],
                        replicas=[r], model_kind=hostapi.MODEL_VIT, max_batch_size=8, pool_size=2,
                        n_workspaces=2) as srv:
        outs = [np.zeros((1, 10), np.float32) for _ in range(5)]
        res = srv.wait(srv.submit_batch([(i, 1, [x[i:i + 1].copy()], [outs[i]]) for i in range(5)]))
        assert all(q.ok == 1 and q.batch_rows == 5 for q in res), [q.error for q in res]
        assert (np.concatenate(outs, 0) == direct).all()


def test_gemm_chain_hook_times_the_gemms_and_leaves_the_replica_usable(tiny):
    """spu_vit_gemm_chain_ms (bench.py's roofline of config E) runs the GEMMs alone on whatever the workspace holds;
    the next real forward is unaffected."""
    vo, w, r = tiny
    x = vo.make_inputs(vo.VIT_TINY, 4, 11)
    before = _run(r, x, 10)
    s = torch.cuda.Stream()
    out = torch.empty(4, 10, device="cuda")
    ms = r.gemm_chain_ms(torch.from_numpy(x).cuda(), out, reps=2, stream=s)
    assert 0.0 < ms < 50.0
    with pytest.raises(Exception):
        r.gemm_chain_ms(torch.from_numpy(x).cuda(), out, reps=2, stream=None)  # the default stream cannot capture
    assert np.array_equal(_run(r, x, 10), before)


def test_vit_auto_graph_replays_and_follows_the_buffers(spu):
    """SPU_FORWARD_AUTO_GRAPH on a ViT replica: eager on the first sighting of a batch size, captured on the second,
    afterwards one graph per batch size whose pixel / logit pointers are patched per launch; bit-equal to eager."""
    from oracle import vit_oracle as vo
    from starpu_inference_server_b200.vit import VitReplica

    w = vo.make_weights(vo.VIT_TINY, 0)
    plain = VitReplica(w, max_batch=6, n_workspaces=1)
    auto = VitReplica(w, max_batch=6, n_workspaces=1, flags=spu.SPU_PREC_BF16 | spu.SPU_FORWARD_AUTO_GRAPH)
    s = torch.cuda.Stream()
    xs = [torch.from_numpy(vo.make_inputs(vo.VIT_TINY, 6, 20 + k)).cuda() for k in range(3)]
    outs = [torch.full((6, 10), float("nan"), device="cuda") for _ in range(3)]
    ref = []
    for x in xs:
        o = torch.empty(6, 10, device="cuda")
        plain.forward_async(x, o, stream=s)
        ref.append(o)
    for rnd in range(3):
        for k in range(3):
            outs[k].fill_(float("nan"))
            auto.forward_async(xs[k], outs[k], stream=s)
            s.synchronize()
            assert torch.equal(outs[k], ref[k]), (rnd, k)
    o4 = torch.empty(4, 10, device="cuda")  # another batch size: its own eager pass, then its own graph
    r4 = torch.empty(4, 10, device="cuda")
    plain.forward_async(xs[0][:4], r4, stream=s)
    for _ in range(3):
        auto.forward_async(xs[0][:4], o4, stream=s)
        s.synchronize()
        assert torch.equal(o4, r4)
    plain.close()
    auto.close()
