"""CPU oracle for the R2D2 learner hot path -- TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import it, and there only as the checker or as the timed CPU baseline.
The product path (``pytorch_r2d2_b200``) never imports this package and fails loudly
when its CUDA library is missing.

Parity status: PINNED.  Every function here was checked, in the build container, against
the unmodified reference imported from ``/root/reference`` (see ``oracle/make_golden.py``);
the outputs of that run are committed as fixtures under ``tests/golden/`` and re-checked
against the oracle by the ``-m "not gpu"`` tests.
"""
