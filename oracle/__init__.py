"""CPU oracle for the MelGAN / MB-MelGAN / FB-MelGAN vocoder inference path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``tts_b200/`` may import this package.
It is imported by ``tests/``, by ``__graft_entry__.smoke()`` (as the checker) and
by ``bench.py`` for the ``cpu_baseline`` / ``--impl reference`` legs (as the thing
timed on the host cores).  The product path is the CUDA library behind
``include/tts_b200.h`` and fails loudly when that library is missing.

Parity status: PINNED.  ``oracle/gen_golden.py`` runs the *unmodified reference*
(imported read-only from /root/reference in the build container) and commits its
inputs/outputs under ``tests/golden/``; ``tests/test_oracle_golden.py`` checks the
restatements in this package against those vectors.  The reference itself ships no
golden vectors or known-answer tests for this path.
"""
