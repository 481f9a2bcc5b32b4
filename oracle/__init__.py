"""CPU oracle for the MEEM hot path — TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is product code.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs may import it,
and there only as the checker / reported baseline, never as the thing measured
or shipped.  The product (``openflash_b200``) never imports this package and
fails loudly when its CUDA library is missing.
"""
