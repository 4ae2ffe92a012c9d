"""numpy restatement of the reference's per-candidate DSM fit (``superdsm/dsm.py``, ``superdsm/objects.py``).

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  PINNED against the reference itself: every function
below is compared with the output of the reference's own code on the same inputs (golden vectors in
``tests/golden/``, made by ``oracle/make_golden.py``), except the interior-point iterations, which live in
``oracle/cpsolver.py`` (parity unpinned, see there).

Each function cites the reference lines it follows (paths relative to ``/root/reference``).
"""
import math

import numpy as np
import scipy.ndimage as ndi
import scipy.sparse

from . import cpsolver

UINT32_MAX = 4294967295


# ---------------------------------------------------------------------------------------------------------------
# a1: region of a candidate   (objects.py:76-128, image.py:75-87)

def cvxprog_region_mask(y_model, y_mask, atoms, footprint, background_margin):
    """omega = y_mask AND isin(atoms, footprint) AND (EDT(y <= 0) <= background_margin)."""
    mask = np.isin(atoms, list(footprint))                                   # objects.py:93
    mask = np.logical_and(y_mask, mask)                                      # image.py:82
    return np.logical_and(mask, ndi.distance_transform_edt(y_model <= 0) <= background_margin)  # objects.py:127


# ---------------------------------------------------------------------------------------------------------------
# a2: normalised pixel coordinates   (image.py:6-21)

def pixel_map(shape, normalized=True):
    z = (np.array(shape) - 1. if normalized else np.ones(2))[..., None, None]
    z[z == 0] = 1
    return np.indices(shape) / z


# ---------------------------------------------------------------------------------------------------------------
# a3: the sub-sampled smooth matrix G~   (dsm.py:137-235), closed form

def gaussian_weights(sigma, radius):
    """The 1-D weights ``scipy.ndimage.gaussian_filter`` convolves with (``_gaussian_kernel1d``)."""
    x = np.arange(-radius, radius + 1)
    phi = np.exp(-0.5 / (sigma * sigma) * x ** 2)
    return phi / phi.sum()


def gaussian_psf(sigma, shape_multiplier):
    """dsm.py:137-142: gaussian_filter of a centred delta on a box of round(1 + 4 sigma mult) pixels.

    scipy truncates its own kernel at int(4 sigma + 0.5); with 'reflect' boundaries and a single centred
    non-zero sample the result is the outer product of the truncated 1-D weights, zero-padded if the box is
    larger than the kernel.
    """
    size = round(1 + sigma * 4 * shape_multiplier)
    c = size // 2
    lw = int(4.0 * float(sigma) + 0.5)
    w = gaussian_weights(sigma, lw)
    line = np.zeros(size)
    for k in range(-lw, lw + 1):
        # reflect ('symmetric' in numpy terms) boundary handling of scipy.ndimage
        i = c + k
        period = 2 * size
        i %= period
        if i >= size: i = period - 1 - i
        line[i] += w[k + lw]
    return np.outer(line, line)


def create_subsample_grid(mask, subsample):
    """dsm.py:164-181 on the bounding-box crop of omega.

    Grid = omega AND lattice (rows, cols = 0 mod subsample), then greedily add the first (row-major) omega
    pixel whose chessboard distance to the grid is minimal among those >= subsample, until none is left.
    """
    grid = np.zeros_like(mask)
    grid[::subsample, ::subsample] = True
    grid &= mask
    rr, cc = np.nonzero(mask)
    gr, gc = np.nonzero(grid)
    if len(gr) > 0:
        dist = np.full(len(rr), UINT32_MAX, dtype=np.int64)
        for k0 in range(0, len(gr), 256):
            dd = np.maximum(np.abs(rr[:, None] - gr[None, k0:k0 + 256]), np.abs(cc[:, None] - gc[None, k0:k0 + 256]))
            dist = np.minimum(dist, dd.min(axis=1))
    else:
        dist = np.full(len(rr), UINT32_MAX, dtype=np.int64)   # distance_transform_bf without any background
    while True:
        outside = dist >= subsample
        if not outside.any(): break
        dmin = dist[outside].min()
        k = int(np.nonzero(dist == dmin)[0][0])                # first in row-major order
        grid[rr[k], cc[k]] = True
        dist = np.minimum(dist, np.maximum(np.abs(rr - rr[k]), np.abs(cc - cc[k])))
    return grid


def numpy_pairwise_sum_f32(a):
    """Pure-Python emulation of numpy's float32 pairwise summation (what ``M.sum(axis=1)`` does, dsm.py:192)."""
    f = np.float32
    n = len(a)
    if n < 8:
        r = f(0.)
        for v in a: r = f(r + v)
        return r
    if n <= 128:
        r = [f(a[k]) for k in range(8)]
        i = 8
        while i < n - (n % 8):
            for k in range(8): r[k] = f(r[k] + a[i + k])
            i += 8
        res = f(f(f(r[0] + r[1]) + f(r[2] + r[3])) + f(f(r[4] + r[5]) + f(r[6] + r[7])))
        while i < n:
            res = f(res + a[i])
            i += 1
        return res
    n2 = n // 2
    n2 -= n2 % 8
    return f(numpy_pairwise_sum_f32(a[:n2]) + numpy_pairwise_sum_f32(a[n2:]))


def smooth_matrix(mask, smooth_amount, shape_multiplier, smooth_subsample, dtype='float32', return_grid=False):
    """dsm.py:184-235.  ``mask`` is the full-frame omega; returns the n x m CSR matrix G~ (float64 container)."""
    n = int(mask.sum())
    null = scipy.sparse.csr_matrix((n, 0), dtype=np.float64)
    if not (smooth_amount < np.inf):
        return (null, None) if return_grid else null
    psf = gaussian_psf(smooth_amount, shape_multiplier).astype(dtype)
    assert psf.shape[0] % 2 == 1, psf.shape[0]                               # dsm.py:147
    R = psf.shape[0] // 2
    rows = np.nonzero(mask.any(axis=1))[0]
    cols = np.nonzero(mask.any(axis=0))[0]
    crop = mask[rows[0]:rows[-1] + 1, cols[0]:cols[-1] + 1] if n > 0 else mask[:0, :0]
    # NB dsm.py:185-186 drops empty rows/columns *inside* the bounding box as well
    crop = mask[rows, :][:, cols] if n > 0 else crop
    if crop.shape[0] <= R or crop.shape[1] <= R:                             # dsm.py:187
        return (null, None) if return_grid else null
    grid = create_subsample_grid(crop, smooth_subsample)
    rr, cc = np.nonzero(crop)
    gr, gc = np.nonzero(grid)
    m = len(gr)
    # nearest-lattice search window: only grid points with |dr|, |dc| <= R are non-zero
    order = np.argsort(gc, kind='stable')
    data, indices, indptr = [], [], [0]
    M_rows = []
    # assemble row by row in blocks to bound memory; row sums are float32 pairwise sums over all m columns
    sums = np.empty(n, dtype=dtype)
    blk = max(1, (1 << 22) // max(m, 1))
    vals_all, cols_all, cnt_all = [], [], np.empty(n, dtype=np.int64)
    for i0 in range(0, n, blk):
        dr = gr[None, :] - rr[i0:i0 + blk, None]
        dc = gc[None, :] - cc[i0:i0 + blk, None]
        inside = (np.abs(dr) <= R) & (np.abs(dc) <= R)
        M = np.zeros(dr.shape, dtype=dtype)
        M[inside] = psf[R + dr[inside], R + dc[inside]]
        s = M.sum(axis=1)                                                    # dsm.py:192
        M /= s[:, None]                                                      # dsm.py:193
        sums[i0:i0 + blk] = s
        M_rows.append(scipy.sparse.csr_matrix(M))
    assert (sums > 0).all()                                                  # dsm.py:194 (row sums)
    mat = scipy.sparse.vstack(M_rows, format='csr').astype(np.float64)
    assert (np.asarray(mat.sum(axis=0)).reshape(-1) > 0).all()               # dsm.py:194 (column sums)
    if return_grid:
        return mat, {'grid_r': gr, 'grid_c': gc, 'pix_r': rr, 'pix_c': cc, 'row_sums': sums, 'R': R,
                     'bbox_rows': rows, 'bbox_cols': cols}
    return mat


# ---------------------------------------------------------------------------------------------------------------
# a4-a7: energy, gradient, Hessian   (dsm.py:48-63, 80-94, 242-385)

class Energy(object):
    """psi(theta, xi) over the pixels of omega.  ``model`` is the full-frame image y, ``mask`` the full-frame omega."""

    def __init__(self, model, mask, epsilon, alpha, smooth_mat, sparsity_tol=0, hessian_sparsity_tol=0):
        assert epsilon > 0 and alpha >= 0
        self.model, self.mask = model, mask
        self.smooth_mat = smooth_mat
        self.x = pixel_map(model.shape)[:, mask]                             # dsm.py:272, image.py:97-103
        self.y = model[mask]                                                 # dsm.py:274
        self.epsilon, self.alpha = epsilon, alpha
        self.sparsity_tol, self.hessian_sparsity_tol = sparsity_tol, hessian_sparsity_tol
        x1, x2 = self.x
        self.q = np.stack([x1 * x1, x2 * x2, 2 * x1 * x2, 2 * x1, 2 * x2, np.ones_like(x1)])  # dsm.py:242-250
        self.m = smooth_mat.shape[1]
        self.n = len(self.y)

    def surface(self, p):
        """dsm.py:86-94 (note the factor 2 on b)."""
        a1, a2, a3, b1, b2, c = p[:6]
        x1, x2 = self.x
        s = (x1 * a1 + x2 * a3) * x1 + (x1 * a3 + x2 * a2) * x2 + 2 * (x1 * b1 + x2 * b2) + c
        if self.m > 0: s = s + self.smooth_mat @ p[6:]
        return s

    def _maps(self, p):
        t = self.y * self.surface(p)
        valid = t >= -np.log(np.finfo(float).max)                            # dsm.py:298
        h = np.full(t.shape, np.nan)
        h[valid] = np.exp(-t[valid])
        theta = np.ones_like(t)
        theta[valid] = h[valid] / (1 + h[valid])                             # dsm.py:306-310
        return t, h, valid, theta

    def __call__(self, p):
        p = np.asarray(p, dtype=float).reshape(-1)
        assert not np.isnan(p).any()                                         # dsm.py:72
        t, h, valid, _ = self._maps(p)
        phi = np.where(valid, np.log(1 + np.where(valid, h, 0)), -t)         # dsm.py:319-322
        value = phi.sum()
        if self.m > 0:
            xi = p[6:]
            o2 = self.alpha * np.sqrt(xi * xi + self.epsilon).sum() - self.alpha * math.sqrt(self.epsilon) * self.m
            value += max(o2, 0)                                              # dsm.py:324-331
        return float(value)

    def grad(self, p):
        p = np.asarray(p, dtype=float).reshape(-1)
        assert not np.isnan(p).any()
        _, _, _, theta = self._maps(p)
        term1 = -self.y * theta                                              # dsm.py:344
        g = self.q @ term1                                                   # dsm.py:345
        if self.m > 0:
            xi = p[6:]
            u = term1.copy()
            u[np.abs(u) < self.sparsity_tol] = 0                             # dsm.py:346
            g2 = self.smooth_mat.T @ u + self.alpha * (xi / np.sqrt(xi * xi + self.epsilon))  # dsm.py:349-350
            g = np.concatenate([g, g2])
        return g

    def hessian(self, p):
        """Full symmetric dense Hessian; the reference returns its block-lower-triangular form (dsm.py:369-371)."""
        p = np.asarray(p, dtype=float).reshape(-1)
        assert not np.isnan(p).any()
        _, _, _, theta = self._maps(p)
        kappa = theta - np.square(theta)                                     # dsm.py:362
        kappa[kappa < self.sparsity_tol] = 0
        d = kappa * self.y * self.y
        Hqq = (self.q * d) @ self.q.T
        if self.m == 0: return Hqq
        G = self.smooth_mat
        Gd = G.multiply(d[:, None]).tocsr()
        Hxq = np.asarray((Gd.T @ self.q.T))
        Hxx = np.asarray((G.T @ Gd).todense())
        xi = p[6:]
        t3 = xi * xi
        t2 = np.sqrt(t3 + self.epsilon)
        gd = self.alpha * (1 / t2 - t3 / np.power(t2, 3))                     # dsm.py:372
        gd[gd < 0] = 0
        H = np.block([[Hqq, Hxq.T], [Hxq, Hxx]])
        H[np.arange(6, 6 + self.m), np.arange(6, 6 + self.m)] += gd
        if self.hessian_sparsity_tol > 0:                                    # dsm.py:376-382: the diagonal is kept
            small = np.abs(H) < self.hessian_sparsity_tol
            small[np.diag_indices_from(small)] = False
            H[small] = 0
        return H


# ---------------------------------------------------------------------------------------------------------------
# a8-a10: cvxprog   (dsm.py:431-490, objects.py:287-412)

class CvxprogError(Exception):
    pass


def _solve(J, x0, scale, trace=None, events=None, restore_policy=None):
    """dsm.CP(...).solve(): cvxopt.solvers.cp on scale * J  (dsm.py:446-490)."""
    def F(x):
        return scale * J(x), scale * J.grad(x)

    def FH(x, z):
        return scale * J(x), scale * J.grad(x), z * scale * J.hessian(x)
    return cpsolver.cp(F, FH, x0, trace=trace, events=events, restore_policy=restore_policy)


def estimate_initialization(model, mask):
    """objects.py:287-296 -> parameter vector of the data ellipse (dsm.py:96-111 with U = I)."""
    fg = np.where(mask, model, 0) > 0
    xmap = pixel_map(model.shape)
    ctr = np.round(ndi.center_of_mass(fg)).astype(int)
    ctr = xmap[:, ctr[0], ctr[1]]
    half = (xmap[:, fg] - ctr[:, None]).std(axis=1)
    half = np.max([half, np.full(half.shape, 1e-8)], axis=0)
    A = np.diag(1. / np.square(half))
    b = A.dot(ctr)
    c = np.inner(ctr, b) - 1
    return np.array([-A[0, 0], -A[1, 1], -A[0, 1], b[0], b[1], -c])


def elliptical_solution(J_ell, scale, info=None):
    """objects.py:321-358."""
    sol, arr, val = None, None, np.inf
    try:
        sol = _solve(J_ell, np.zeros(6), scale)
        arr = sol['x']
        val = J_ell(arr)
    except Exception:
        pass
    if info is not None: info['pass1'] = None if sol is None else (sol['status'], sol['iterations'])
    if sol is None or sol['status'] != 'optimal':
        init = estimate_initialization(J_ell.model, J_ell.mask)
        init_val = J_ell(init)
        if not (init_val > val):
            try:
                sol = _solve(J_ell, init, scale)
                arr = sol['x']
                val = J_ell(arr)
                if info is not None: info['pass2'] = (sol['status'], sol['iterations'])
            except Exception as err:
                if sol is None: raise CvxprogError(err)
    assert arr is not None
    return arr


def cvxprog(model, mask, scale, epsilon, alpha, smooth_amount, smooth_subsample, gaussian_shape_multiplier,
            smooth_mat_dtype='float32', sparsity_tol=0, hessian_sparsity_tol=0, init='elliptical', info=None, **unused):
    """objects.py:361-412.  Returns (J, params, status)."""
    G = smooth_matrix(mask, smooth_amount, gaussian_shape_multiplier, smooth_subsample, smooth_mat_dtype)
    J = Energy(model, mask, epsilon, alpha, G, sparsity_tol, hessian_sparsity_tol)
    scale_cp = scale / G.shape[0]
    if callable(init):
        params = init(G.shape[1])
    else:
        if init == 'elliptical':
            J_ell = Energy(model, mask, epsilon, alpha, scipy.sparse.csr_matrix((G.shape[0], 0)))
            params = elliptical_solution(J_ell, scale_cp, info)
        else:
            params = np.zeros(6)
        params = np.concatenate([params, np.zeros(G.shape[1])])
    try:
        trace = [] if info is not None else None
        events = [] if info is not None else None
        if info is not None: info['dsm_trace'], info['dsm_events'], info['init'] = trace, events, np.asarray(params, dtype=float).copy()
        sol = _solve(J, params, scale_cp, trace=trace, events=events, restore_policy=unused.get('restore_policy'))
        solution = sol['x']
        if info is not None: info['dsm'] = (sol['status'], sol['iterations'])
        if sol['status'] == 'unknown' and J(solution) > J(params):
            status = 'fallback'
        else:
            status = 'optimal'
    except Exception as err:
        if info is not None: info['dsm_error'] = f'{type(err).__name__}: {err}'
        status = 'fallback'
    if status == 'fallback': solution = params
    return J, np.asarray(solution, dtype=float).reshape(-1), status


# ---------------------------------------------------------------------------------------------------------------
# a11: result of one candidate   (objects.py:148-212, dsm.py:113-128, _aux.py:15-23)

def extract_foreground_fragment(fg_mask):
    if fg_mask.any():
        rows, cols = fg_mask.any(axis=1), fg_mask.any(axis=0)
        rmin, rmax = np.where(rows)[0][[0, -1]]
        cmin, cmax = np.where(cols)[0][[0, -1]]
        return np.array([rmin, cmin]), fg_mask[rmin:rmax + 1, cmin:cmax + 1]
    return np.zeros(2, int), np.zeros((1, 1), bool)


def map_to_image_pixels(p, shape, pad):
    """dsm.py:113-128 for a region that spans the whole image (offset (0, 0), roi shape = image shape)."""
    A = np.array([[p[0], p[2]], [p[2], p[1]]])
    b = np.asarray(p[3:5], dtype=float)
    c = p[5]
    roi_max = np.array(shape) - 1.
    G = np.diag(1. / roi_max)
    v = -G.dot(np.add((0, 0), pad))
    A2 = G.dot(A).dot(G)
    b2 = G.dot(A.dot(v) + b)
    c2 = np.inner(v, A.dot(v)) + 2 * np.inner(b, v) + c
    return A2, b2, c2


def compute_object(y_model, y_mask, atoms, footprint, dsm_cfg, info=None):
    """objects.py:177-212.  Returns a dict with the fields the reference stores on the ``Object``."""
    cfg = dict(dsm_cfg)
    cfg.pop('smooth_mat_max_allocations', None)
    mask = cvxprog_region_mask(y_model, y_mask, atoms, footprint, cfg.pop('background_margin'))
    if (y_model[mask] > 0).sum() == 1:                                       # objects.py:184-191
        return dict(fg_offset=np.zeros(2, int), fg_fragment=np.zeros((1, 1), bool), energy=0., on_boundary=False,
                    is_optimal=False, fallback=False, params=None, mask=mask)
    J, p, status = cvxprog(y_model, mask, info=info, **cfg)
    A2, b2, c2 = map_to_image_pixels(p, y_model.shape, pad=1)
    xm = pixel_map(np.add(y_model.shape, 2), normalized=False).reshape(2, -1)
    s = np.einsum('ij,ij->i', np.dot(xm.T, A2), xm.T) + 2 * np.inner(xm.T, b2) + c2
    s = s.reshape(np.add(y_model.shape, 2))
    if J.m > 0:
        s[1:-1, 1:-1][mask] += J.smooth_mat @ p[6:]
    padded_fg = s > 0
    fg = padded_fg[1:-1, 1:-1]
    if fg.any():
        off, frag = extract_foreground_fragment(np.logical_and(mask, fg))
    else:
        off, frag = np.zeros(2, int), np.zeros((1, 1), bool)
    on_boundary = bool(padded_fg[0].any() or padded_fg[-1].any() or padded_fg[:, 0].any() or padded_fg[:, -1].any())
    return dict(fg_offset=off, fg_fragment=frag, energy=J(p), on_boundary=on_boundary,
                is_optimal=(status == 'optimal'), fallback=(status == 'fallback'), params=p, mask=mask, n=J.n, m=J.m)
