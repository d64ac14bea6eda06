"""CPU oracle for the StarPU inference codelet hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is product code: only
``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl
reference`` legs of ``bench.py`` may import or execute it, and only as the
checker or the timed CPU baseline -- never as the thing shipped.  The product
path (``starpu_inference_server_b200``) fails loudly when its CUDA library is
missing; it never falls back to this package.
"""
