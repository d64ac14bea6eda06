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
