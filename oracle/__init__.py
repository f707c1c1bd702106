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


def warm_up():
    """Run one tiny forward of the network oracle and throw it away.

    Measured on the GPU boxes' hosts (tools/cpu_oracle_first_call.py): in about 1 process of 20
    the FIRST torch-CPU forward of the process differs from every later identical call by ~3.5e-5
    (the later calls agree with float64 to 5e-7); after any earlier forward it never did (0 of
    40).  The checker must not be the noisy side of a 1e-5 comparison, so everything that uses
    the oracle as a checker calls this first."""
    import torch
    from . import net_oracle as NO
    with torch.no_grad():
        p = NO.init_params(1, 64)
        NO.qnet_forward(p, torch.rand(8, 4, 84, 84), None, None)
