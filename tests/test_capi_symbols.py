"""CPU: the C-ABI library builds for sm_100a, loads, and exports every symbol that
include/spu_b200.h declares.  No compute calls (there is no GPU here) -- except that the
create call must FAIL loudly without an sm_100 device (no CPU fallback exists)."""
import ctypes as C
import os
import re

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "spu_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(spu_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_all_exported(built_libs):
    from starpu_inference_server_b200 import capi

    L = capi.lib()
    declared = _declared_symbols()
    assert len(declared) >= 17
    for name in declared:
        assert hasattr(L, name), f"{name} declared in spu_b200.h but not exported"
    assert sorted(capi.SYMBOLS) == declared
    assert L.spu_abi_version() == 2


def test_library_is_sm100a_only(built_libs):
    import subprocess

    out = subprocess.run(["/usr/local/cuda/bin/cuobjdump", "-lelf", built_libs[0]], capture_output=True,
                         text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_no_cpu_fallback(built_libs):
    """Without a usable sm_100 device every create call fails with UNSUPPORTED."""
    import torch

    if torch.cuda.is_available():
        import pytest

        pytest.skip("GPU present")
    from starpu_inference_server_b200 import capi

    L = capi.lib()
    assert L.spu_device_count() == 0
    desc = capi.BertDesc(1, 64, 1, 64, 10, 8, 2, 1e-12)
    n = 5 + 16
    bufs = [np.zeros(64 * 64, np.float32) for _ in range(n)]
    arr = (C.c_void_p * n)(*[b.ctypes.data for b in bufs])
    h = C.c_void_p()
    st = L.spu_bert_replica_create(C.byref(desc), arr, n, 0, 1, 8, 1, 0, C.byref(h))
    assert st == capi.SPU_ERR_UNSUPPORTED and not h.value
    assert b"no CPU fallback" in L.spu_last_error()
