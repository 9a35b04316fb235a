"""CPU oracle for the mixing-style DSP front end.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the shipped
product: only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import it, and there only as the
checker or as the timed CPU baseline.  The product path
(``mixing_style_transfer_b200``) never imports this package and has no CPU
fallback.

Parity status: PINNED.  The restatement in ``mixfe_oracle.py`` is checked
(bit-for-bit where the same torch kernels are reached, else to 1e-6) against the
reference's own modules imported from ``/root/reference/src`` in the build
container (``oracle/validate_against_reference.py``) and against the golden
vectors that script family wrote to ``tests/golden/`` (``tests/golden/make_golden.py``).
"""
