"""CPU oracle for the GP-MPC hot path -- TEST INFRASTRUCTURE ONLY.

This package is a numpy/scipy *restatement* of the reference algorithms
(shiivashaakeri/gp-mpc-rocket-landing, mounted read-only at /root/reference in the
build container) for the per-control-step numeric core: GP Gram / fit / predict,
feature extraction, horizon uncertainty propagation, OSQP-form QP assembly, an
OSQP-style ADMM solve and the GP-MPC / RTI orchestration.  Every function cites
the reference file:line it follows.

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import this package, and only as the
checker or as the timed CPU baseline.  The product (``gp_mpc_rocket_landing_b200``)
never imports it and has no CPU fallback.

Parity pinning (see DESIGN.md "Oracle"):

* GP path (kernels, exact/sparse/fast GP, features, Simple3DoFGP/StructuredRocketGP):
  PINNED -- ``oracle/gen_golden.py`` imports the *real* reference ``src.gp`` (through
  a ``sys.modules`` shim that bypasses the ``simdyn`` import) and writes
  ``tests/golden/*.npz``; ``tests/test_oracle_golden.py`` checks this restatement
  against those outputs.
* QP assembly (``src/mpc/osqp_rti.py`` P, q, A, l, u): PINNED the same way (the
  reference assembler run with a stub ``osqp`` module and the Euler dynamics below).
* Uncertainty propagation: PINNED via the reference ``UncertaintyPropagator`` run
  on a 14-state embedding of the 7-state model.
* QP *solve* (OSQP, third-party, ``osqp>=0.6.0``, not installed, no source on disk),
  dynamics (``simdyn``, undeclared third-party, absent) and ``GPMPC.solve``
  (CasADi/IPOPT, absent): **parity unpinned** -- the reference holds no tests or
  vectors for them; the restatements follow the published OSQP algorithm and the
  in-repo Euler model and are cross-checked by an independent KKT solve.
"""
