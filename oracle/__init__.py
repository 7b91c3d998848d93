"""CPU oracles for the pilottai_b200 hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is product code: only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it, and only as the checker or the timed
CPU baseline.  The product path (``pilottai_b200``) never imports this package
and fails loudly when its CUDA library is missing.
"""
