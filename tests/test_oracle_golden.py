"""The oracle against the golden vectors produced by the reference's own code (tests/golden/, oracle/make_golden.py)
and against the reference's own unit-test fixtures.  CPU only."""
import numpy as np
import pytest
import scipy.sparse

from oracle import dsm_oracle as orc
from oracle import cpsolver
from tests.golden_util import CASES, GoldenCase, probe_matrix


@pytest.fixture(scope='module', params=CASES)
def case(request):
    return GoldenCase(request.param)


def test_region_mask_reference_fixture():
    """tests/test_objects.py:50-74 of the reference."""
    y = np.array([[-1, -1, -1, -1, -1], [-1, -1, -1, -1, -1], [-1, -1, -1, -1, -1],
                  [-1, +1, -1, -1, -1], [-1, +1, -1, -1, +1], [-1, +1, -1, -1, +1]])
    atoms = np.array([[1, 1, 1, 1, 1], [1, 1, 1, 1, 1], [1, 1, 1, 1, 2], [1, 1, 1, 2, 2], [1, 1, 1, 2, 2], [1, 1, 1, 2, 2]])
    expected = np.array([[0, 0, 0, 0, 0], [0, 1, 0, 0, 0], [1, 1, 1, 0, 0], [1, 1, 1, 0, 0], [1, 1, 1, 0, 0], [1, 1, 1, 0, 0]], bool)
    got = orc.cvxprog_region_mask(y, np.ones(y.shape, bool), atoms, {1}, 2)
    assert np.array_equal(got, expected)


def test_pixel_map_reference_fixture():
    """tests/test_image.py of the reference: get_pixel_map((6, 3), normalized=True)."""
    m = orc.pixel_map((6, 3))
    assert np.allclose(m[0][:, 0], np.arange(6) / 5.) and np.allclose(m[1][0], np.arange(3) / 2.)
    assert np.array_equal(orc.pixel_map((1, 4))[0], np.zeros((1, 4)))       # zero-size guard (image.py:20)


def test_gaussian_psf_matches_scipy():
    import scipy.ndimage as ndi
    for sigma, mult, exact in ((4, 2, True), (10, 2, True), (8, 2, True), (5, 2, True), (2.5, 2, True), (1.5, 2, True), (7, 1, False)):
        size = round(1 + sigma * 4 * mult)
        inp = np.zeros((size, size))
        inp[size // 2, size // 2] = 1
        ref = ndi.gaussian_filter(inp, sigma)
        got = orc.gaussian_psf(sigma, mult)
        # box >= scipy's own kernel (the configurations the reference ships): one product per entry, bit-equal;
        # box smaller than the kernel: reflected terms add up, equal to rounding
        assert np.array_equal(got, ref) if exact else np.allclose(got, ref, rtol=1e-14, atol=0)


def test_pairwise_sum_emulation_matches_numpy():
    rng = np.random.default_rng(0)
    for m in list(range(1, 20)) + [63, 64, 65, 127, 128, 129, 255, 257, 300, 777]:
        M = rng.random((20, m)).astype(np.float32)
        M[rng.random((20, m)) < 0.7] = 0
        assert np.array_equal(M.sum(axis=1), np.array([orc.numpy_pairwise_sum_f32(r) for r in M], np.float32))


def test_region_masks(case):
    for k, fp in enumerate(case.footprints):
        got = orc.cvxprog_region_mask(case.y, case.y_mask, case.atoms, fp, case.dsm_cfg['background_margin'])
        assert np.array_equal(got, case.mask(k)), (case.name, k)


def test_smooth_matrix_bit_exact(case):
    cfg = case.dsm_cfg
    for k in range(len(case)):
        G = orc.smooth_matrix(case.mask(k), cfg['smooth_amount'], cfg['gaussian_shape_multiplier'], cfg['smooth_subsample'])
        G.sort_indices()
        ref = case.G(k)
        assert G.shape == ref.shape, (case.name, k)
        assert np.array_equal(G.indptr, ref.indptr) and np.array_equal(G.indices, ref.indices), (case.name, k)
        assert np.array_equal(G.data, ref.data), (case.name, k)


def test_energy_grad_hessian(case):
    cfg = case.dsm_cfg
    for k in range(len(case)):
        J = orc.Energy(case.y, case.mask(k), cfg['epsilon'], cfg['alpha'], case.G(k))
        for pt in case.energy_points(k):
            p = pt['p']
            assert J(p) == pytest.approx(pt['value'], rel=1e-12, abs=1e-12)
            g = J.grad(p)
            assert np.allclose(g, pt['grad'], rtol=1e-11, atol=1e-11 * max(1, np.abs(pt['grad']).max()))
            H = J.hessian(p)
            assert np.allclose(H, H.T, rtol=0, atol=1e-12 * max(1, np.abs(H).max()))
            probe = H @ probe_matrix(len(p))
            assert np.allclose(probe, pt['hess_probe'], rtol=1e-10, atol=1e-10 * max(1, np.abs(probe).max()))
            if pt['hess_lower'] is not None:
                assert np.allclose(np.tril(H), pt['hess_lower'], rtol=1e-10, atol=1e-12 * max(1, np.abs(H).max()))


def test_compute_object(case):
    for k, fp in enumerate(case.footprints):
        res = orc.compute_object(case.y, case.y_mask, case.atoms, fp, case.dsm_cfg)
        ref = case.fit(k)
        assert res['energy'] == pytest.approx(float(ref['energy']), rel=1e-9, abs=1e-9), (case.name, k)
        assert bool(res['is_optimal']) == bool(ref['is_optimal']) and bool(res['fallback']) == bool(ref['fallback'])
        assert bool(res['on_boundary']) == bool(ref['on_boundary'])
        assert np.array_equal(res['fg_offset'], ref['fg_offset']) and np.array_equal(res['fg_fragment'], ref['fg_fragment'])


def test_solver_certificate():
    """The restated cp solver on a problem with a known optimum: logistic regression optimum has zero gradient."""
    case = GoldenCase('hand')
    cfg = case.dsm_cfg
    J = orc.Energy(case.y, case.mask(0), cfg['epsilon'], cfg['alpha'], scipy.sparse.csr_matrix((int(case.mask(0).sum()), 0)))
    sol = orc._solve(J, np.zeros(6), cfg['scale'] / J.n)
    assert sol['status'] == 'optimal' and sol['iterations'] < 40
    assert np.abs(J.grad(sol['x'])).max() < 1e-6 * J.n
    # a singular Hessian at the start point must raise cvxopt's rank error (all-equal pixels -> rank-1 Hessian)
    def F(x): return 0.0, np.zeros(3)
    def FH(x, z): return 0.0, np.zeros(3), np.zeros((3, 3))
    with pytest.raises(ValueError, match='Rank'):
        cpsolver.cp(F, FH, np.zeros(3))


def test_first_pass_trials_do_not_use_F(case, monkeypatch):
    """What the CUDA solver relies on when it runs first-pass line searches without evaluations (csrc/ctrl.cuh PC_TRIALREQ):
    while relaxed steps are allowed, cp decides a first-pass trial on newgap alone and never uses what F returned there
    (cpsolver.py: the `i == 0` branch of the line search).  Replacing those values by NaN changes no solve, bit for bit;
    and the line searches of the two passes of an iteration examine the same number of trials (the predictor of the
    speculative requests) in the overwhelming majority of iterations."""
    fits = []
    for poison in (False, True):
        monkeypatch.setattr(cpsolver, 'POISON_UNUSED', poison)
        monkeypatch.setattr(cpsolver, 'LS_LOG', [])
        fits.append([orc.compute_object(case.y, case.y_mask, case.atoms, fp, case.dsm_cfg) for fp in case.footprints])
        log = list(cpsolver.LS_LOG)
    for a, b in zip(*fits):
        assert a['energy'] == b['energy'] and a['fallback'] == b['fallback'] and a['is_optimal'] == b['is_optimal']
        if a['params'] is not None: assert np.array_equal(np.asarray(a['params']), np.asarray(b['params']))
        assert np.array_equal(a['fg_fragment'], b['fg_fragment'])
    pairs = [(p, q) for p, q in zip(log[:-1], log[1:]) if p[1] == 0 and q[1] == 1 and p[0] == q[0]]
    assert len(pairs) > 20
    assert np.mean([p[3] == q[3] for p, q in pairs]) > 0.9
    # a first pass that accepted its first trial is followed by a second pass that does, too (the trial evaluated with its Hessian)
    assert all(q[3] == 1 for p, q in pairs if p[3] == 1)
