"""CPU oracle for the NFOPP optimisation inner loop.  TEST INFRASTRUCTURE ONLY.

This package is a CPU restatement (numpy float64 / torch float32, analytic gradients, no autograd) of the
reference's hot path.  It exists to *check* the CUDA product path; it is never a fallback for it.
Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` leg
may import it.  The product package (``pytorch-motion-planner_b200/nfopp_b200``) must not.

Parity status: **pinned by live execution of the unmodified reference** (the reference's own tests hold no
numeric golden vectors).  ``oracle/make_golden.py`` runs the reference modules from
``/root/reference`` in this container, and the vectors it produced are committed under ``tests/golden/``;
``tests/test_oracle_golden.py`` checks every oracle function against them on every run (no reference needed),
and ``tests/test_oracle_vs_reference.py`` re-checks against the live reference whenever it is present.

Every function cites the reference ``file:line`` it follows (paths relative to the reference root).
"""
