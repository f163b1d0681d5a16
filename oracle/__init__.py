"""CPU oracle for the ArcNerf volumetric-rendering hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
legs may import this package, and only as the checker or the timed CPU baseline.  Nothing under
``arcnerf_b200/`` imports it; the product path fails loudly when its CUDA library is missing.
"""
