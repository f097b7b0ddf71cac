"""CPU oracle for the NLBAC neural-ODE hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import this package.  The product
path (the ``nlbac_b200`` package) never imports it and has no CPU fallback.

Parity status: **pinned against the reference's own Python** (the classes in
``/root/reference/*/sac_cbf_clf/model.py`` imported in the build container,
``tests/golden/make_golden.py``) for the NODE models, ``train_step`` and the
MLPs; **unpinned upstream** for the solver itself, because ``torchdiffeq 0.2.3``
(README.md:33 of the reference) is a third-party dependency that is neither
vendored under /root/reference nor installable here.  ``torchdiffeq_restated``
restates its published algorithm and is anchored by analytic known answers
(matrix exponential, convergence orders, the rk4 3/8-rule closed form).
"""
