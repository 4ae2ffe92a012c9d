"""Round-2 rows of the hot-path scope table on the device (through the C ABI): several images per launch, the level-synchronous
c2f scheduler, cvxprog on a single-foreground-pixel region, plan-carrying handles.  -m gpu."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def eng():
    from superdsm_b200.objects import get_engine
    return get_engine(0)


def _toy(seed):
    from workloads import synth
    return synth.make_workload('toy', oversegment=2, max_set_size=3, seed=seed)


def test_multi_image_batch_is_bitwise_the_single_image_result(eng):
    """Candidates of three images in ONE launch (sdsm_set_images / sdsm_cand.img): every result bit-identical to the image solved
    alone -- energies, parameters, statuses, foreground masks."""
    ws = [_toy(s) for s in (7, 8, 9)]
    cfg = ws[0]['dsm_cfg']
    singles = []
    for w in ws:
        eng.set_image(w['y'], w['y_mask'], w['atoms'], cfg['background_margin'])
        eng.set_params(cfg)
        singles.append(eng.solve(w['footprints']))
    handles = [eng.upload_image(w['y'], w['y_mask'], w['atoms'], cfg['background_margin']) for w in ws]
    eng.bind_images(handles)
    eng.set_params(cfg)
    res = eng.solve_multi([w['footprints'] for w in ws])
    assert len(res) == sum(len(w['footprints']) for w in ws)
    k = 0
    for i, (w, single) in enumerate(zip(ws, singles)):
        for j in range(len(single)):
            assert int(res.cands['img'][k]) == i
            assert res.status[k] == single.status[j] and res.energy[k] == single.energy[j], (i, j)
            assert np.array_equal(res.params(k), single.params(j))
            (o1, f1), (o2, f2) = res.foreground(k), single.foreground(j)
            assert np.array_equal(o1, o2) and np.array_equal(f1, f2)
            k += 1
    # an empty image in the set is allowed, and the seam-level helper sets the callers' objects in place
    from superdsm_b200.objects import compute_objects_multi
    from tests.compat.refshim import Image, Object
    jobs = [([Object(fp) for fp in w['footprints'][:5]], Image(w['y'], w['y_mask']), w['atoms']) for w in ws]
    jobs[1] = ([], jobs[1][1], jobs[1][2])
    fallbacks = compute_objects_multi(jobs, cfg)
    assert len(fallbacks) == 3 and fallbacks[1] == 0
    for j, ob in enumerate(jobs[2][0]): assert ob.energy == singles[2].energy[j]


def test_prepared_handles_stay_valid_across_plans(eng):
    """A handle of Engine.prepare carries its own plan (the workspace layout is a function of the plan, not of hidden context
    state): planning another batch in between does not disturb it."""
    import torch
    w = _toy(7)
    eng.set_image(w['y'], w['y_mask'], w['atoms'], w['dsm_cfg']['background_margin'])
    eng.set_params(w['dsm_cfg'])
    h1 = eng.prepare(w['footprints'][:10])
    eng.run_prepared(h1); torch.cuda.synchronize()
    e1 = h1['out']['energy'][:80].cpu().numpy().view(np.float64).copy()
    h2 = eng.prepare(w['footprints'][3:])           # replans
    eng.run_prepared(h2); torch.cuda.synchronize()
    eng.run_prepared(h1); torch.cuda.synchronize()
    assert np.array_equal(h1['out']['energy'][:80].cpu().numpy().view(np.float64), e1)
    st = eng.solver_stats(h1['ws'], h1['plan'])
    assert st['done'] == 10 and not st['abort']
    assert np.array_equal(h1['out']['solver'][:16].cpu().numpy().view(np.uint64), [10, 0])


def test_cvxprog_fits_a_region_with_one_foreground_pixel(eng):
    """cvxprog itself fits such a region (objects.py:361-412); only _compute_object skips it (objects.py:184-191)."""
    from oracle import dsm_oracle as orc
    from superdsm_b200.objects import cvxprog, compute_objects
    from tests.compat.refshim import Image, Object
    rng = np.random.default_rng(3)
    y = -0.2 + 0.05 * rng.standard_normal((40, 48))
    y[20, 24] = 0.4                                   # the only positive pixel
    mask = np.zeros(y.shape, bool); mask[8:32, 10:40] = True
    assert (y[mask] > 0).sum() == 1
    kw = dict(scale=1000, epsilon=1.0, alpha=0.5, smooth_amount=np.inf, smooth_subsample=20, gaussian_shape_multiplier=2,
              smooth_mat_dtype='float32', init='elliptical')
    J, model, status = cvxprog(Image(y, mask), smooth_mat_allocation_lock=None, **kw)
    Jo, po, so = orc.cvxprog(y, mask, **kw)
    assert status == so
    assert J(model) == pytest.approx(Jo(po), rel=1e-6, abs=2e-7 * mask.sum() / 1000)
    # ... while the batch seam skips it: energy 0, not optimal
    ob = Object({1})
    cfg = dict(kw, background_margin=1e9, smooth_amount=4, smooth_subsample=8, cachesize=1, cachetest=None, sparsity_tol=0, cp_timeout=300)
    compute_objects([ob], Image(y, np.ones(y.shape, bool)), mask.astype(np.int32), cfg, None)
    assert ob.energy == 0.0 and ob.is_optimal is False


def test_c2f_level_synchronous_scheduler(eng):
    """c2freganal.py:193-287 with every cluster's split state machine advanced in lock-step: the regions of a round (<= 2 per
    live cluster) go through ONE batch with no deformation parameters; the image is uploaded once.  The normalised energy of
    every region any split attempt asked for equals the oracle's fit of that region (1e-6) wherever the fit is regular."""
    from oracle import dsm_oracle as orc
    from superdsm_b200.c2f import NormalizedEnergyComputer, run_level_synchronous
    from tests.compat.c2f_cluster import cluster_task
    from tests.compat.refshim import Image
    import scipy.ndimage as ndi
    from workloads import synth
    w = synth.make_workload('toy', oversegment=1, max_set_size=1, seed=21, shape=(200, 240), n=9)
    y = Image(w['y'], np.ones(w['y'].shape, bool))
    comp, ncomp = ndi.label(w['y'] > 0)
    sizes = ndi.sum(w['y'] > 0, comp, index=np.arange(1, ncomp + 1))
    keep = [k + 1 for k in range(ncomp) if sizes[k] >= 80]
    markers = np.where(np.isin(comp, keep), comp, 0)
    idx = ndi.distance_transform_edt(markers == 0, return_distances=False, return_indices=True)
    clusters = markers[idx[0], idx[1]]
    cfg = dict(w['dsm_cfg'])
    args = dict(dsm_cfg=cfg, max_atom_norm_energy=0.05, min_atom_radius=4, min_norm_energy_improvement=0.1)

    logs = {lab: [] for lab in keep}
    computer = NormalizedEnergyComputer(w['y'], cfg, eng)
    out = run_level_synchronous([cluster_task(y, clusters, lab, y.mask, log=logs[lab], **args) for lab in keep], computer)
    n_requests = sum(1 if e[0] == 'root' else 2 for lg in logs.values() for e in lg)
    assert computer.uploads == 1                                       # one upload of y per image
    assert computer.launches <= run_level_synchronous.last_rounds < n_requests      # batched across the clusters (a round of cache hits launches nothing)
    assert len(out) == len(keep) and any(len(lg) > 1 for lg in logs.values()), 'no cluster was split: the test image is too easy'

    # the oracle on every region the state machines asked for (each region once)
    keys = ('scale', 'epsilon', 'alpha', 'smooth_subsample', 'gaussian_shape_multiplier', 'smooth_mat_dtype', 'init')
    worst, irregular, compared = 0.0, 0, 0
    for key, (rq, bits, iters) in computer.solved.items():
        full = np.zeros(w['y'].shape, bool)
        (r, c), (h, ww) = rq.offset, rq.mask.shape
        full[r:r + h, c:c + ww] = rq.mask
        info = {}
        J, p, _ = orc.cvxprog(w['y'], full, smooth_amount=np.inf, info=info, **{k: cfg[k] for k in keys})
        ref, got = J(p) / full.sum(), computer.cache[key]
        if not isinstance(got, float): continue
        err = abs(got - ref) / max(abs(ref), 1e-300)
        # a fit that left the regular regime on either side (an elliptical pass or the final solve ending 'unknown' / at the
        # iteration limit, a failed factorisation) may stop elsewhere: counted, and bounded below
        regular = bits == 0 and iters.max() < 100 and info.get('pass1', ('', 0))[0] == 'optimal' and info.get('dsm', ('', 0))[0] == 'optimal' \
            and not info.get('dsm_events') and max([abs(t['relaxed']) for t in info.get('dsm_trace') or [{'relaxed': 0}]]) <= 1
        compared += 1
        if err <= 1e-6 or abs(got - ref) <= 2e-10: worst = max(worst, err); continue
        assert not regular, (rq.offset, rq.mask.sum(), got, ref, bits, iters, info.get('pass1'), info.get('dsm'))
        irregular += 1
    assert compared >= 50 and irregular <= 0.05 * compared, (irregular, compared)
    print('clusters', len(keep), 'requests', n_requests, 'rounds', run_level_synchronous.last_rounds, 'regions', compared, 'irregular mismatches', irregular, 'worst relative error', worst)


@pytest.mark.parametrize('name,kw', [('toy', dict()), ('bbbc039', dict()), ('gowt1', dict(sigma2=42.43)), ('nih3t3', dict(lower_clip_mean=True)),
                                     ('toy', dict(offset_clip=np.inf, sigma2=7.5)), ('tiled4k', dict(sigma2=10, shape=(777, 1031)))])
def test_preprocessing_on_the_device_is_bit_exact(eng, name, kw):
    """preprocess.py:39-68 (three Gaussian filters, sigma up to 42: 341 taps with 'reflect' boundaries, the EDT blend) against the scipy
    statement of the same lines: identical bits on the five synthetic shapes."""
    import scipy.ndimage as ndi
    from superdsm_b200.preprocess import offset_image
    from workloads import synth
    kw = dict(kw)
    spec = dict(synth.WORKLOADS[name])
    if 'shape' in kw: spec['shape'] = kw.pop('shape'); spec['n'] = 300
    g_raw, _, _ = synth.make_image(**spec)
    sigma1, sigma2 = kw.get('sigma1', np.sqrt(2)), kw.get('sigma2', 40)
    offset_clip, lower = kw.get('offset_clip', 3), kw.get('lower_clip_mean', False)
    # the reference's lines, verbatim in meaning (scipy on the host)
    offset = ndi.gaussian_filter(g_raw, sigma2)
    if not np.isinf(offset_clip):
        clip_abs = offset_clip * g_raw.std()
        clipped = ndi.gaussian_filter(g_raw.clip(0, clip_abs), sigma2)
        t = ndi.distance_transform_edt(~(g_raw > clip_abs))
        t = (sigma2 - t).clip(0, np.inf)
        t = (t / t.max()) ** 2
        offset = (1 - t) * clipped + t * offset
    if lower: offset = np.max([offset, np.full(g_raw.shape, g_raw.mean())], axis=0)
    ref = ndi.gaussian_filter(g_raw, sigma1) - offset
    y = offset_image(g_raw, sigma1, sigma2, offset_clip, lower, engine=eng)
    assert y.shape == ref.shape and np.array_equal(y, ref), (np.abs(y - ref).max(), int((y != ref).sum()))


def test_hessian_sparsity_tol_against_the_reference(eng):
    """dsm/hessian_sparsity_tol (dsm.py:376-382): off-diagonal entries of the Hessian of psi below the tolerance are dropped, the
    diagonal is kept.  Golden: the reference's own Energy.hessian (tests/golden/hesstol.npz, oracle/make_hesstol_golden.py)."""
    import os
    from tests.golden_util import GOLDEN_DIR, GoldenCase
    z = np.load(os.path.join(GOLDEN_DIR, 'hesstol.npz'))
    for name, ks in (('hand', (0, 2)), ('toy', (0, 3))):
        case = GoldenCase(name)
        eng.set_image(case.y, case.y_mask, case.atoms, case.dsm_cfg['background_margin'])
        for k in ks:
            tol, p, ref = float(z[f'{name}_{k}_tol']), z[f'{name}_{k}_p'], z[f'{name}_{k}_hess_lower']
            eng.set_params(dict(case.dsm_cfg, hessian_sparsity_tol=tol))
            eng.solve(case.footprints)
            _, _, H = eng.eval_candidate(k, p)
            L = np.tril(H)
            assert np.array_equal(L != 0, ref != 0), (name, k, int((L != 0).sum()), int((ref != 0).sum()))     # the same entries survive
            assert np.allclose(L, ref, rtol=1e-9, atol=1e-12 * np.abs(ref).max())
            eng.set_params(case.dsm_cfg)
            eng.solve(case.footprints)
            assert (np.tril(eng.eval_candidate(k, p)[2]) != 0).sum() > (ref != 0).sum()                         # without it: denser


def test_speculative_evaluation_changes_no_bit(monkeypatch):
    """The solver evaluates line-search trials ahead of cp's decisions (several backtracking points per request, the Hessian
    with a trial that will be accepted, no evaluations at all in first-pass searches that do not look at F): decisions and
    results must be those of the sequential search, bit for bit -- energies, parameters, status, iteration and evaluation
    counters, foreground masks.  Context A is created with every speculation off, context B with the defaults."""
    from superdsm_b200.engine import Engine
    from workloads import synth
    for var in ('SDSM_SPEC_MAX', 'SDSM_SPEC_HESS', 'SDSM_SPEC_SKIP', 'SDSM_SPEC_SOLO', 'SDSM_SPEC_PRIO'): monkeypatch.delenv(var, raising=False)
    monkeypatch.setenv('SDSM_SPEC_MAX', '1'); monkeypatch.setenv('SDSM_SPEC_HESS', '0'); monkeypatch.setenv('SDSM_SPEC_SKIP', '0')
    plain = Engine(0)
    for var in ('SDSM_SPEC_MAX', 'SDSM_SPEC_HESS', 'SDSM_SPEC_SKIP'): monkeypatch.delenv(var)
    spec = Engine(0)
    monkeypatch.setenv('SDSM_SPEC_PRIO', '0'); monkeypatch.setenv('SDSM_SPEC_RETRY', '8')
    eager = Engine(0)                                                    # every candidate speculates, as far as it goes
    w = synth.make_workload('gowt1', oversegment=3, max_set_size=4, seed=1)
    fps = w['footprints'][::3]                                           # every size class, incl. the long irregular chains
    out = []
    for e in (plain, spec, eager):
        e.set_image(w['y'], w['y_mask'], w['atoms'], w['dsm_cfg']['background_margin'])
        e.set_params(w['dsm_cfg'])
        res = e.solve(fps)
        out.append((res.energy.copy(), res.status.copy(), res.info.copy(), [res.params(k) for k in range(len(fps)) if res.status[k] < 2],
                    res.fgbox.copy(), res.fgmask.copy(), e.solver_stats()))
    a = out[0]
    assert (a[2][:, 6] > 40).any() and (a[1] == 0).mean() > 0.9
    for b in out[1:]:
        assert np.array_equal(a[0].view(np.uint64), b[0].view(np.uint64))
        assert np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2]) and np.array_equal(a[4], b[4]) and np.array_equal(a[5], b[5])
        for p, q in zip(a[3], b[3]): assert np.array_equal(p.view(np.uint64), q.view(np.uint64))
    # and it does what it is for: fewer control tasks (round trips) and fewer evaluation items
    assert out[1][6]['n_ctrl'] < 0.7 * a[6]['n_ctrl'] and out[1][6]['n_eval'] < 0.9 * a[6]['n_eval']
