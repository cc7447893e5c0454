"""CPU oracle for the spectral token-mixer hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import this package.  Nothing under
``vit_spectre_experiments_b200/`` imports it: the product path fails loudly
when the CUDA library is missing instead of falling back to this code.
"""
