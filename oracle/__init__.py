"""CPU oracle for the synthesis hot path (HiFi-GAN V1 Generator + LengthRegulator).

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product: only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it, and there only as the checker or as the
timed CPU baseline.  The product path (``visual_onoma_to_wave_b200``) never imports it
and fails loudly when the CUDA library is missing.

Parity status: the reference ships no tests and no golden vectors for this path
("parity unpinned" by the reference itself).  The oracle is
pinned by differential execution: ``tests/golden/make_golden.py`` imports the
reference's own modules from /root/reference in the build container, runs them on
seeded inputs, and commits inputs/outputs as fixtures under ``tests/golden/``; the
``-m "not gpu"`` tests check every oracle function against those fixtures.
"""
