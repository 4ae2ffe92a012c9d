"""CPU oracle for the SuperDSM per-candidate DSM convex solve.

TEST INFRASTRUCTURE ONLY.  Nothing under ``superdsm_b200/`` may import this
package; only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` use it, and only as the checker / the
timed CPU baseline -- never as the product path.

Parity status
-------------
* ``dsm_oracle`` (region mask, sub-sample grid, smooth matrix G~, energy, gradient,
  Hessian, result rasterisation) is PINNED: it is checked against golden vectors
  produced by importing the reference's own ``superdsm/dsm.py``, ``objects.py`` and
  ``image.py`` unchanged (``oracle/ref_harness.py`` + ``oracle/make_golden.py``,
  fixtures under ``tests/golden/``), and against the reference's own unit-test
  fixtures (``tests/test_objects.py:50-74``, ``tests/test_image.py``).
* ``cpsolver`` (the Newton/KKT interior-point iterations) is a restatement of
  ``cvxopt.solvers.cp`` (cvxopt 1.2.7, pinned in the reference's ``requirements.txt:7``
  and ``superdsm.yml``).  cvxopt is an un-vendored third-party dependency that is absent
  from ``/root/reference`` and cannot be installed offline, and the reference holds no
  golden vector at that boundary.  PARITY UNPINNED for the
  iteration arithmetic; what is checked instead are optimality certificates
  (gradient norm at the returned point, energy <= any other converged solver's energy).
"""
