"""Restatement of ``cvxopt.solvers.cp`` for the one use SuperDSM makes of it.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  PARITY UNPINNED: cvxopt 1.2.7
(``/root/reference/requirements.txt:7``) is an absent third-party dependency; this file
restates its published algorithm (``cvxopt/cvxprog.py``: ``cp`` -> ``cpl``;
``cvxopt/misc.py``: ``kkt_chol2``, ``compute_scaling``, ``update_scaling``) specialised
to the problem the reference hands it at ``superdsm/dsm.py:488``:

    minimize f(x)           (no G, h, A, b; cvxopt options left at their defaults)

``cp`` rewrites this in epigraph form  min t  s.t.  f(x) - t <= 0  and calls ``cpl`` with
c = (0, 1), one nonlinear constraint, scalar slack s, multiplier z and Nesterov-Todd
scaling d = sqrt(s/z), lambda = sqrt(s z).  With no linear constraints the ``chol2`` KKT
solver reduces to a Cholesky factorisation of  H = z * Hess f(x).

The CUDA solver (``superdsm_b200/csrc/ctrl.cuh``) implements exactly the state machine
below; the two are compared iterate-for-iterate in ``tests/``.
"""
import math

import numpy as np

STEP = 0.99
BETA = 0.5
ALPHA = 0.01
EXPON = 3
MAX_RELAXED_ITERS = 8
MAXITERS = 100
ABSTOL = 1e-7
RELTOL = 1e-6
FEASTOL = 1e-7
REFINEMENT = 1

RANK_ERROR = 'Rank(A) < p or Rank([H(x); A; Df(x); G]) < n'

# Policy for a factorisation failure while a saved line-search state exists.  cvxopt restores the saved state only for
# 0 < relaxed_iters < MAX_RELAXED_ITERS and gives up ('unknown') when the failure happens right after the last relaxed
# step (relaxed_iters == MAX_RELAXED_ITERS).  Whether the factorisation of such a matrix fails at all is decided by the
# last bits of a numerically singular z * Hess f far away from the optimum (LAPACK, CHOLMOD, numpy and the CUDA solver
# disagree there), and when it happens to succeed cp restores the saved state one step later anyway (newphi >= phi0).
# With 'saved' (the policy the CUDA solver implements) the saved state is restored in both cases, so
# that the outcome no longer depends on those bits; 'cvxopt' is the strict restatement.  Second difference: after a restore
# cvxopt refines the Newton system of the restored point against the gradient and Hessian of the ABANDONED point (its residual
# closures are not updated) -- garbage by construction, and when the excursion overflowed a NaN that turns the whole solve
# into an exception; 'saved' uses the derivatives of the restored point there.
RESTORE_POLICY = 'saved'

#: diagnostics: a list here collects (iteration, pass, first step, trials, accepted step) of every line search
LS_LOG = None

#: test hook.  cp evaluates F at every trial point of the FIRST pass of an iteration, but while relaxed steps are allowed
#: (0 <= relaxed_iters < MAX_RELAXED_ITERS) the trial is accepted or rejected on newgap alone -- a function of the step -- and
#: nothing F returned there is used afterwards.  With POISON_UNUSED the values F returns at those trials are replaced by NaN:
#: a solve that is unchanged by it proves the claim on that solve (tests/test_oracle_golden.py); the CUDA solver does not make
#: those evaluations (csrc/ctrl.cuh PC_TRIALREQ).
POISON_UNUSED = False


class _State(object):
    __slots__ = ('x', 't', 's', 'z', 'd', 'lmbda', 'rx_x', 'rx_t', 'rznl', 'dx', 'dt', 'dz', 'ds')


def _chol_lower(H):
    """Cholesky factor of the symmetric matrix whose LOWER triangle is stored in H.

    Raises ArithmeticError like ``lapack.potrf`` / ``cholmod.numeric`` do for a matrix that
    is not (numerically) positive definite.
    """
    Hl = np.tril(H)
    Hs = Hl + np.tril(Hl, -1).T
    if not np.isfinite(Hs).all():
        raise ArithmeticError('non-finite KKT matrix')
    try:
        return np.linalg.cholesky(Hs)
    except np.linalg.LinAlgError as err:
        raise ArithmeticError(str(err))


def _chol_solve(L, b):
    from scipy.linalg import solve_triangular
    return solve_triangular(L, solve_triangular(L, b, lower=True), lower=True, trans='T')


def _symv_lower(H, u):
    Hl = np.tril(H)
    return Hl @ u + np.tril(Hl, -1).T @ u


def cp(F, FH, x0, trace=None, maxiters=MAXITERS, restore_policy=None, events=None):
    """Minimise f.

    :param F: ``F(x) -> (f, g)``, value and gradient (1-D array) of f at x.
    :param FH: ``FH(x, z) -> (f, g, H)`` where ``H = z * Hess f(x)``; only the lower triangle
        of H is used (the reference passes a block-lower-triangular sparse matrix,
        ``dsm.py:369-371,470-471``).
    :param x0: start point.
    :param trace: optional list; one dict per iteration is appended (used by the tests to compare
        the CUDA solver's iterates).
    :param restore_policy: ``'saved'`` or ``'cvxopt'`` (see ``RESTORE_POLICY`` above); default: the module's policy.
    :param events: optional list; ``('chol_fail', iteration, relaxed_iters)`` is appended for every failed factorisation.
    :return: dict with the keys the reference reads (``objects.py:300,329,397-399``).
    :raises ValueError: with cvxopt's rank message if the first factorisation fails.
    """
    if restore_policy is None: restore_policy = RESTORE_POLICY
    assert restore_policy in ('saved', 'cvxopt')
    restore_upto = MAX_RELAXED_ITERS + 1 if restore_policy == 'saved' else MAX_RELAXED_ITERS
    x = np.array(x0, dtype=float).reshape(-1).copy()
    t = 0.0
    s = 1.0
    z = 1.0
    relaxed_iters = 0
    saved = None
    d = lmbda = None
    theta1 = theta2 = theta3 = None
    resx0 = resznl0 = pres0 = dres0 = None
    phi0 = dphi0 = gap0 = step0 = dsdz0 = sigma0 = eta0 = None

    def result(status, gap, relgap, pcost, dcost, pres, dres):
        return {'status': status, 'x': x.copy(), 'gap': gap, 'relative gap': relgap,
                'primal objective': pcost, 'dual objective': dcost,
                'primal slack': s, 'dual slack': z,
                'primal infeasibility': pres, 'dual infeasibility': dres, 'iterations': iters}

    for iters in range(maxiters + 1):
        f0, g, H = FH(x, z)
        g = np.asarray(g, dtype=float).reshape(-1)
        f = f0 - t

        gap = s * z
        rx_x = z * g
        rx_t = 1.0 - z
        resx = math.sqrt(float(rx_x @ rx_x) + rx_t * rx_t)
        rznl = s + f
        resznl = abs(rznl)

        pcost = t
        dcost = pcost + z * rznl - gap
        if pcost < 0.0:
            relgap = gap / -pcost
        elif dcost > 0.0:
            relgap = gap / dcost
        else:
            relgap = None
        pres = math.sqrt(resznl ** 2)
        dres = resx
        if iters == 0:
            resx0 = max(1.0, resx)
            resznl0 = max(1.0, resznl)
            pres0 = max(1.0, pres)
            dres0 = max(1.0, dres)
            gap0 = gap
            theta1 = 1.0 / gap0
            theta2 = 1.0 / resx0
            theta3 = 1.0 / resznl0
        phi = theta1 * gap + theta2 * resx + theta3 * resznl
        pres = pres / pres0
        dres = dres / dres0

        if trace is not None:
            trace.append({'iter': iters, 'x': x.copy(), 't': t, 's': s, 'z': z, 'f': f0, 'gap': gap,
                          'pres': pres, 'dres': dres, 'relaxed': relaxed_iters})

        if (pres <= FEASTOL and dres <= FEASTOL and
                (gap <= ABSTOL or (relgap is not None and relgap <= RELTOL))) or iters == maxiters:
            status = 'unknown' if iters == maxiters else 'optimal'
            return result(status, gap, relgap, pcost, dcost, pres, dres)

        if iters == 0:
            d = math.sqrt(s / z)
            lmbda = math.sqrt(s * z)
        lmbdasq = lmbda * lmbda

        gk = g  # gradient used by cp's KKT wrapper and by cpl's residual closure (fDf)
        Hk = H  # Hessian used by cpl's residual closure (fH)
        try:
            L = _chol_lower(H)
        except ArithmeticError:
            singular_kkt_matrix = False
            if events is not None: events.append(('chol_fail', iters, relaxed_iters))
            if iters == 0:
                raise ValueError(RANK_ERROR)
            elif 0 < relaxed_iters < restore_upto:
                # Restart a standard line search from the last saved state.
                phi, gap = phi0, gap0
                d, lmbda = saved.d, saved.lmbda
                x = saved.x.copy()
                t = saved.t
                s, z = saved.s, saved.z
                lmbdasq = lmbda * lmbda
                rx_x, rx_t = saved.rx_x.copy(), saved.rx_t
                resx = math.sqrt(float(rx_x @ rx_x) + rx_t * rx_t)
                rznl = saved.rznl
                resznl = abs(rznl)
                relaxed_iters = -1
                try:
                    # cp's wrapper re-evaluates F at the restored point for the factorisation and for
                    # gradf0; cpl's own Df/H closures keep the values of the abandoned point.
                    _, g_r, H_r = FH(x, z)
                    L = _chol_lower(H_r)
                    gk = np.asarray(g_r, dtype=float).reshape(-1)
                    if restore_policy == 'saved':
                        # ... which may be garbage or non-finite: the 'saved' policy refines the Newton system against the
                        # derivatives of the point it restored (g and Hk feed the residual closure below)
                        g, Hk = gk, H_r
                except ArithmeticError:
                    singular_kkt_matrix = True
            else:
                singular_kkt_matrix = True
            if singular_kkt_matrix:
                return result('unknown', gap, relgap, pcost, dcost, pres, dres)

        g_res, H_res = g, Hk

        def f3(bx, bt, bz):
            # cp's epigraph wrapper around kkt_chol2 (no G, no A): returns ux, ut and the SCALED uz.
            a = bz
            ux = _chol_solve(L, bx + gk * bt)
            uz = -d * bt
            ut = float(gk @ ux) + d * d * bt - a
            return ux, ut, uz

        def f4_no_ir(bx, bt, bz, bs):
            bs = bs / lmbda
            bz = bz - d * bs
            ux, ut, uz = f3(bx, bt, bz)
            us = bs - uz
            return ux, ut, uz, us

        def res(ux, ut, uz, us, vx, vt, vz, vs):
            # residual of the Newton equations (cpl's ``res``), using cpl's own Df and H
            vx = vx - _symv_lower(H_res, ux)
            wz3 = uz / d
            vx = vx - g_res * wz3
            vt = vt + wz3
            vz = vz - (float(g_res @ ux) - ut)
            vz = vz - d * us
            vs = vs - lmbda * (uz + us)
            return vx, vt, vz, vs

        def f4(bx, bt, bz, bs):
            ux, ut, uz, us = f4_no_ir(bx, bt, bz, bs)
            for _ in range(REFINEMENT):
                wx, wt, wz, ws = res(ux, ut, uz, us, bx, bt, bz, bs)
                ex, et, ez, es = f4_no_ir(wx, wt, wz, ws)
                ux = ux + ex
                ut += et
                uz += ez
                us += es
            return ux, ut, uz, us

        sigma, eta = 0.0, 0.0
        for i in (0, 1):
            mu = gap
            ds = -lmbdasq + sigma * mu
            dx = -(1.0 - eta) * rx_x
            dt = -(1.0 - eta) * rx_t
            dz = -(1.0 - eta) * rznl
            dx, dt, dz, ds = f4(dx, dt, dz, ds)

            dsdz = ds * dz
            dz2 = dz / d
            ds2 = ds * d
            ds = ds / lmbda
            dz = dz / lmbda
            tmax = max(0.0, -ds, -dz)
            step = 1.0 if tmax == 0 else min(1.0, STEP / tmax)
            ls_first, ls_trials = step, 0

            # "Backtrack until newx is in domain of f": F never leaves its domain here, but the call is
            # made (and cached by the reference, dsm.py:388-418).
            F(x + step * dx)

            phi = theta1 * gap + theta2 * resx + theta3 * resznl
            if i == 0:
                dphi = -phi
            else:
                dphi = -theta1 * (1 - sigma) * gap - theta2 * (1 - eta) * resx - theta3 * (1 - eta) * resznl

            backtrack = True
            while backtrack:
                newx = x + step * dx
                newt = t + step * dt
                newz = z + step * dz2
                news = s + step * ds2
                nf0, ng = F(newx)
                ls_trials += 1
                ng = np.asarray(ng, dtype=float).reshape(-1)
                if POISON_UNUSED and i == 0 and 0 <= relaxed_iters < MAX_RELAXED_ITERS: nf0, ng = float('nan'), np.full_like(ng, np.nan)
                newf = nf0 - newt
                newrx_x = newz * ng
                newrx_t = 1.0 - newz
                newresx = math.sqrt(float(newrx_x @ newrx_x) + newrx_t * newrx_t)
                newrznl = news + newf
                newresznl = abs(newrznl)
                newgap = (1.0 - (1.0 - sigma) * step) * gap + step ** 2 * dsdz
                newphi = theta1 * newgap + theta2 * newresx + theta3 * newresznl

                if i == 0:
                    if newgap <= (1.0 - ALPHA * step) * gap and \
                            (0 <= relaxed_iters < MAX_RELAXED_ITERS or newphi <= phi + ALPHA * step * dphi):
                        backtrack = False
                        sigma = min(newgap / gap, (newgap / gap) ** EXPON)
                        eta = 0.0
                    else:
                        step *= BETA
                else:
                    if relaxed_iters == -1 or (relaxed_iters == 0 == MAX_RELAXED_ITERS):
                        if newphi <= phi + ALPHA * step * dphi:
                            # cvxopt has ``relaxed_iters == 0`` here (a comparison, not an assignment)
                            backtrack = False
                        else:
                            step *= BETA
                    elif relaxed_iters == 0 < MAX_RELAXED_ITERS:
                        if newphi <= phi + ALPHA * step * dphi:
                            relaxed_iters = 0
                        else:
                            phi0, dphi0, gap0 = phi, dphi, gap
                            step0 = step
                            saved = _State()
                            saved.d, saved.lmbda = d, lmbda
                            saved.x, saved.t = x.copy(), t
                            saved.s, saved.z = s, z
                            saved.rx_x, saved.rx_t, saved.rznl = rx_x.copy(), rx_t, rznl
                            saved.dx, saved.dt = dx.copy(), dt
                            saved.dz, saved.ds = dz, ds          # already divided by lambda
                            dsdz0 = dsdz
                            sigma0, eta0 = sigma, eta
                            relaxed_iters = 1
                        backtrack = False
                    elif 0 <= relaxed_iters < MAX_RELAXED_ITERS > 0:
                        if newphi <= phi0 + ALPHA * step0 * dphi0:
                            relaxed_iters = 0
                        else:
                            relaxed_iters += 1
                        backtrack = False
                    elif relaxed_iters == MAX_RELAXED_ITERS > 0:
                        if newphi <= phi0 + ALPHA * step0 * dphi0:
                            backtrack = False
                            relaxed_iters = 0
                        elif newphi >= phi0:
                            # Resume the last saved line search.
                            phi, dphi, gap = phi0, dphi0, gap0
                            step = step0
                            d, lmbda = saved.d, saved.lmbda
                            lmbdasq = lmbda * lmbda
                            x, t = saved.x.copy(), saved.t
                            s, z = saved.s, saved.z
                            rx_x, rx_t, rznl = saved.rx_x.copy(), saved.rx_t, saved.rznl
                            dx, dt = saved.dx.copy(), saved.dt
                            dz, ds = saved.dz, saved.ds
                            # cvxopt restores dz, ds (lambda-scaled) but not dz2, ds2; with the scaling W
                            # restored, the unscaled steps of the saved search are recomputed identically:
                            dz2 = dz * lmbda / d
                            ds2 = ds * lmbda * d
                            dsdz = dsdz0
                            sigma, eta = sigma0, eta0
                            relaxed_iters = -1
                        elif newphi <= phi + ALPHA * step * dphi:
                            backtrack = False
                            relaxed_iters = -1
                        else:
                            # cvxopt falls through all branches and repeats the same trial for ever; the
                            # reference's SIGALRM time-out (dsm.py:483-489) is what ends it.  Treat as the
                            # time-out would: an exception.
                            raise TimeoutError('cpl line search stalled')
            if LS_LOG is not None: LS_LOG.append((iters, i, ls_first, ls_trials, step))

        x = x + step * dx
        t = t + step * dt
        ds = 1.0 + step * ds
        dz = 1.0 + step * dz
        ds = ds * lmbda
        dz = dz * lmbda
        # update_scaling (nonlinear block)
        sq_s = math.sqrt(ds)
        sq_z = math.sqrt(dz)
        d = d * sq_s / sq_z
        lmbda = sq_s * sq_z
        s = lmbda * d
        z = lmbda / d

    raise AssertionError('unreachable')


# ---------------------------------------------------------------------------------------------------------------
# cvxopt-protocol adapter, so that the reference's own ``dsm.CP`` (dsm.py:431-490) can drive this solver through a
# stub ``cvxopt`` module (oracle/ref_harness.py).

class matrix(np.ndarray):
    """Minimal stand-in for ``cvxopt.matrix``: a float ndarray; 1-D input becomes a column."""

    def __new__(cls, data=0.0, size=None, tc='d'):
        arr = np.array(data, dtype=float)
        if arr.ndim == 0:
            arr = np.full(size if size is not None else (1, 1), float(arr))
        elif arr.ndim == 1:
            arr = arr.reshape(-1, 1)
        return arr.view(cls)


class spmatrix(object):
    """Minimal stand-in for ``cvxopt.spmatrix`` (COO triplets; duplicate entries add up)."""

    def __init__(self, vals, rows, cols, size=None, tc='d'):
        import scipy.sparse
        self.size = tuple(size)
        self.coo = scipy.sparse.coo_matrix((np.asarray(vals, float), (np.asarray(rows), np.asarray(cols))), shape=self.size)

    def toarray(self):
        return self.coo.toarray()


def _dense(H):
    return H.toarray() if isinstance(H, spmatrix) else np.asarray(H, dtype=float)


def cp_cvxopt_protocol(Fc):
    """``cvxopt.solvers.cp(F)`` for a callable following cvxopt's F(), F(x), F(x, z) protocol."""
    mnl, x0 = Fc()
    assert mnl == 0
    x0 = np.asarray(x0, dtype=float).reshape(-1)

    def F(x):
        f, Df = Fc(matrix(x))
        return float(np.asarray(f).reshape(-1)[0]), np.asarray(Df, dtype=float).reshape(-1)

    def FH(x, z):
        f, Df, H = Fc(matrix(x), matrix([z]))
        return float(np.asarray(f).reshape(-1)[0]), np.asarray(Df, dtype=float).reshape(-1), _dense(H)

    sol = cp(F, FH, x0)
    sol['x'] = matrix(sol['x'])
    return sol
