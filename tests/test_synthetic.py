"""CPU: the package's seeded generators draw exactly the oracle's streams, so bench.py and
the parity tests talk about the same weights/requests."""
import numpy as np

from oracle import bert_oracle as bo
from starpu_inference_server_b200 import synthetic
from starpu_inference_server_b200.bert import BertConfig


def test_weights_and_requests_match_oracle_streams():
    spec = bo.BERT_TINY
    cfg = BertConfig(**spec.to_dict())
    a, b = bo.make_weights(spec, 3), synthetic.make_weights(cfg, 3)
    assert list(a) == list(b)
    for k in a:
        assert np.array_equal(a[k], b[k]), k
    for min_len in (None, 4):
        i0, m0 = bo.make_inputs(spec, 5, 40, seed=9, min_len=min_len)
        i1, m1 = synthetic.make_requests(cfg, 5, 40, seed=9, min_len=min_len)
        assert np.array_equal(i0, i1) and np.array_equal(m0, m1)


def test_vit_weights_and_images_match_oracle_streams():
    from oracle import vit_oracle as vo
    from starpu_inference_server_b200.vit import VitConfig

    spec = vo.VIT_TINY
    cfg = VitConfig(layers=spec.layers, hidden=spec.hidden, heads=spec.heads, mlp=spec.mlp, patch=spec.patch,
                    image=spec.image, channels=spec.channels, classes=spec.classes, ln_eps=spec.ln_eps)
    a, b = vo.make_weights(spec, 5), synthetic.make_vit_weights(cfg, 5)
    assert sorted(a) == sorted(b)
    for k in a:
        assert np.array_equal(a[k], b[k]), k
    assert np.array_equal(vo.make_inputs(spec, 3, seed=8), synthetic.make_images(cfg, 3, seed=8))
