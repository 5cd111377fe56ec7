"""TEST INFRASTRUCTURE ONLY — CPU oracle for the MoCA-Video hot path.

Nothing under ``oracle/`` is product code.  Only ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may import it, and only as
the checker.  The product (``moca_video_b200``) never imports this package and fails loudly
when its CUDA library is missing.

Parity pinning: the reference ships no golden vectors or tests.  The
restatements here are pinned by fixtures under ``tests/golden/`` that were produced by
importing the *unmodified* reference from ``/root/reference`` (``oracle/make_golden.py``).
"""
