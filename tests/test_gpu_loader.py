"""GPU: replicas created by the C++ TorchScript loader (libspu_loader.so) from artifacts exported the way the
reference exports them give logits bit-equal to replicas created from the same weights as numpy arrays, and the
loader + replica registry path serves requests (per_device and per_worker replication)."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ld(built_libs):
    from starpu_inference_server_b200 import capi, loader

    assert torch.cuda.is_available() and capi.lib().spu_device_is_sm100(0) == 1
    return loader


@pytest.mark.parametrize("spec_name", ["BERT_TINY", "BERT_BASE"])
def test_bert_replica_from_torchscript_equals_replica_from_arrays(ld, tmp_path, spec_name):
    import hf_keyed
    from oracle import bert_oracle as bo
    from starpu_inference_server_b200.bert import BertReplica

    spec = getattr(bo, spec_name)
    w = bo.make_weights(spec, 0)
    path = str(tmp_path / "bert_libtorch.pt")
    hf_keyed.export_bert_torchscript(spec, w, path)
    ids, mask = bo.make_inputs(spec, 4, 96, seed=3, min_len=9)
    ids_d, mask_d = torch.from_numpy(ids).cuda(), torch.from_numpy(mask).cuda()
    with ld.TorchScriptModule(path, ld.KIND_BERT) as m:
        a = m.create_replica(device=0, max_batch=4, max_seq=128, n_workspaces=2)
        b = m.create_replica(device=0, max_batch=4, max_seq=128)  # clone_model_to_gpus: several replicas per module
    ref = BertReplica(w, max_batch=4, max_seq=128)
    outs = [torch.full((4, 96, spec.hidden), float("nan"), device="cuda") for _ in range(3)]
    for r, o in zip((a, b, ref), outs):
        r.forward_async(ids_d, mask_d, o)
    torch.cuda.synchronize()
    assert torch.equal(outs[0], outs[2]) and torch.equal(outs[1], outs[2])
    want = bo.bert_forward(w, spec, ids, mask)
    assert np.abs(outs[0].cpu().numpy() - want).max() / np.abs(want).max() < 2e-2
    for r in (a, b, ref):
        r.close()


def test_vit_replica_from_scripted_torchvision_equals_replica_from_arrays(ld, tmp_path):
    from oracle import vit_oracle as vo
    from starpu_inference_server_b200.vit import VitReplica

    spec = vo.VIT_TINY
    w = vo.make_weights(spec, 0)
    path = str(tmp_path / "vit.pt")
    torch.jit.script(vo.torchvision_model(w, spec)).save(path)  # models/import_vit.py:51-55
    px = torch.from_numpy(vo.make_inputs(spec, 3, seed=2)).cuda()
    with ld.TorchScriptModule(path, ld.KIND_VIT) as m:
        a = m.create_replica(device=0, max_batch=4)
    ref = VitReplica(w, max_batch=4)
    oa, ob = (torch.full((3, spec.classes), float("nan"), device="cuda") for _ in range(2))
    a.forward_async(px, oa)
    ref.forward_async(px, ob)
    torch.cuda.synchronize()
    assert torch.equal(oa, ob)
    want = vo.vit_forward(w, spec, px.cpu().numpy())
    assert np.abs(oa.cpu().numpy() - want).max() / np.abs(want).max() < 2e-2
    a.close()
    ref.close()


def test_loaded_replicas_serve_requests_per_device_and_per_worker(ld, tmp_path):
    """TorchScript file -> loader -> one replica per assignment -> server (ReplicaRegistry) -> task path; with
    gpu_model_replication per_worker every CUDA worker runs its OWN replica (select_gpu_module by worker id)."""
    import hf_keyed
    from oracle import bert_oracle as bo
    from starpu_inference_server_b200 import hostapi as host

    spec = bo.BERT_TINY
    w = bo.make_weights(spec, 1)
    path = str(tmp_path / "bert_libtorch.pt")
    hf_keyed.export_bert_torchscript(spec, w, path)
    S, H = 32, spec.hidden
    ids, mask = bo.make_inputs(spec, 4, S, seed=4, min_len=3)
    want = bo.bert_forward(w, spec, ids, mask)
    io = dict(inputs=[("input_ids", "i64", [1, S]), ("attention_mask", "i64", [1, S])],
              outputs=[("output0", "f32", [1, S, H])], model_kind=host.MODEL_BERT, max_batch_size=4, pool_size=4)
    for per_worker, n_rep in ((False, 1), (True, 2)):
        with ld.TorchScriptModule(path, ld.KIND_BERT) as m:
            reps = [m.create_replica(device=0, max_batch=4, max_seq=S, n_workspaces=2) for _ in range(n_rep)]
        with host.Server(**io, replicas=reps, device_ids=[0], nworker_per_cuda=2, n_workspaces=2,
                         per_worker_replicas=per_worker) as srv:
            for worker in (0, 1):
                outs = [np.zeros((1, S, H), np.float32) for _ in range(4)]
                res = srv.wait(srv.submit_batch([(i, 1, [ids[i:i + 1].copy(), mask[i:i + 1].copy()], [outs[i]])
                                                 for i in range(4)], fixed_worker=worker))
                assert all(r.ok == 1 and r.worker_id == worker for r in res), res[0].error
                got = np.concatenate(outs, 0)
                assert np.abs(got - want).max() / np.abs(want).max() < 2e-2
            # a per_worker server with a replica missing for a worker is refused at creation
        for r in reps:
            r.close()
    with ld.TorchScriptModule(path, ld.KIND_BERT) as m:
        one = m.create_replica(device=0, max_batch=4, max_seq=S, n_workspaces=2)
    with pytest.raises(host.HostError, match="missing replica for device 0 worker 1"):
        host.Server(**io, replicas=[one], device_ids=[0], nworker_per_cuda=2, n_workspaces=2, per_worker_replicas=True)
    one.close()
