"""CPU oracle for the batched controller hot path -- TEST INFRASTRUCTURE ONLY.

Nothing in the product package may import this.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl reference``
legs use it, and only as the checker or the timed CPU arm.

Parity status:
  * linearization, DARE/LQR gain, LQR closed loop: PINNED -- restatement checked
    against the reference's own Python (imported in the build container) and the
    run log the reference ships (G-LQR), committed under tests/golden/.
  * MPC QP: the reference's arithmetic is CVXPY->OSQP, absent from this image and
    from /root/reference (un-vendored, floor-only ``cvxpy>=1.4``).  The restatement
    follows mpc_controller.py line by line into a QP solved by HiGHS (bundled in
    scipy) + an exact KKT polish; it is pinned on the 200 solves of the shipped
    G-MPC run log (tests/golden/g_mpc.npz) to OSQP's own 1e-5-level accuracy.
"""
