"""CPU oracle for the SFNO spectral hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import this package.  Nothing under
``makani_b200/`` imports it; the product path has no CPU fallback.
"""
