"""Generates ``tests/golden/*.npz`` by running the REFERENCE's own code (``/root/reference/superdsm``) on small seeded
inputs.  Run in the build container only:  ``python -m oracle.make_golden``.

What each ``<case>.npz`` pins, per candidate k:
* ``mask_k``      -- ``Object.get_cvxprog_region`` (objects.py:95-128), bit-packed,
* ``G_k_*``       -- ``SmoothMatrixFactory.get`` (dsm.py:217-235): the CSR matrix G~ bit for bit,
* ``E_k_j_*``     -- ``Energy.__call__/grad/hessian`` (dsm.py:312-385) at three parameter points,
* ``fit_k_*``     -- ``objects._compute_object`` (objects.py:177-212) results, i.e. the reference's cvxprog driver
  on top of the RESTATED cp solver (cvxopt itself is absent; these pin everything but the iteration arithmetic).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_harness                     # noqa: E402
from workloads import synth                    # noqa: E402

OUT = os.path.join(ROOT, 'tests', 'golden')


def toy_cases():
    """(name, y, y_mask, atoms, footprints, dsm_cfg)"""
    cases = []
    w = synth.make_workload('toy', oversegment=2, max_set_size=3)
    fps = w['footprints']
    cases.append(('toy', w['y'], w['y_mask'], w['atoms'], fps[:8] + fps[10:12] + fps[-2:], w['dsm_cfg']))
    # hand-set hyper-parameters off the automation lattice (sigma_G 5, sub-sample 11: R = 20 < 2 * 11), larger m
    w2 = synth.make_workload('toy', oversegment=1, max_set_size=2, seed=11, shape=(220, 260), n=5)
    cfg2 = synth.dsm_config(None, smooth_amount=5, smooth_subsample=11, background_margin=12, alpha=0.2)
    cases.append(('odd', w2['y'], w2['y_mask'], w2['atoms'], [w2['footprints'][k] for k in (0, 2, 6)], cfg2))
    return cases


def hand_cases():
    """Hand-made regions for the edge cases: empty rows inside the bounding box, no lattice point inside omega,
    bounding box not larger than the kernel radius (null matrix), a single foreground pixel."""
    rng = np.random.default_rng(5)
    H, W = 70, 90
    yy, xx = np.mgrid[0:H, 0:W]
    y = 0.5 * np.exp(-(((yy - 30) / 14.) ** 2 + ((xx - 40) / 18.) ** 2)) - 0.2 + rng.normal(0, 0.03, (H, W))
    y_mask = np.ones((H, W), bool)
    atoms = np.ones((H, W), np.int32)
    atoms[:, 45:] = 2
    atoms[40:, :] += 2              # labels 1..4
    atoms[36:38, :] = 5             # a band that splits {1,2} from {3,4}: empty rows for the footprint {1,3}
    atoms[0:3, 0:3] = 6             # tiny region -> null matrix
    y[0:3, 0:3] = -0.1
    y[1, 1] = 0.2                   # exactly one foreground pixel in atom 6
    cfg = synth.dsm_config(None, smooth_amount=4, smooth_subsample=8, background_margin=9, alpha=0.05)
    fps = [frozenset([1]), frozenset([1, 3]), frozenset([1, 2, 3, 4]), frozenset([6]), frozenset([2, 5])]
    return [('hand', y, y_mask, atoms, fps, cfg)]


def main():
    ref = ref_harness.load()
    os.makedirs(OUT, exist_ok=True)
    rng = np.random.default_rng(123)
    for name, y, y_mask, atoms, footprints, dsm_cfg in toy_cases() + hand_cases():
        y_img = ref.image.Image.create_from_array(y, normalize=False, mask=y_mask)
        x_map = y_img.get_map(normalized=False, pad=1)
        out = dict(y=y, y_mask=y_mask, atoms=atoms.astype(np.int32),
                   fp_flat=np.concatenate([sorted(fp) for fp in footprints]).astype(np.int32),
                   fp_len=np.array([len(fp) for fp in footprints], np.int32))
        for key, value in dsm_cfg.items():
            if key in ('cachetest', 'init', 'smooth_mat_dtype'): continue
            out['cfg_' + key] = np.float64(value)
        for k, fp in enumerate(footprints):
            obj = ref.objects.Object()
            obj.footprint = set(fp)
            region = obj.get_cvxprog_region(y_img, atoms, dsm_cfg['background_margin'])
            out[f'mask_{k}'] = np.packbits(region.mask)
            # smooth matrix + energy derivatives straight from the reference classes
            factory = ref.dsm.SmoothMatrixFactory(dsm_cfg['smooth_amount'], dsm_cfg['gaussian_shape_multiplier'],
                                                  dsm_cfg['smooth_subsample'], None, dsm_cfg['smooth_mat_dtype'])
            import contextlib, io
            with contextlib.redirect_stdout(io.StringIO()):
                J = ref.dsm.Energy(region, dsm_cfg['epsilon'], dsm_cfg['alpha'], factory)
            G = J.smooth_mat.tocsr()
            G.sort_indices()
            assert np.array_equal(G.data.astype(np.float32).astype(np.float64), G.data)   # fp32 values in an fp64 container
            out[f'G_{k}_data'] = G.data.astype(np.float32)
            out[f'G_{k}_indices'] = G.indices.astype(np.int32)
            out[f'G_{k}_indptr'] = G.indptr.astype(np.int32)
            out[f'G_{k}_shape'] = np.array(G.shape, np.int64)
            m = G.shape[1]
            pts = [np.zeros(6 + m),
                   np.concatenate([[-3., -2., .5, 1.2, .9, -.3], 0.3 * rng.standard_normal(m)]),
                   np.concatenate([[-40., -25., 3., 11., 9., -3.], 2.0 * rng.standard_normal(m)])]
            for j, p in enumerate(pts):
                out[f'E_{k}_{j}_p'] = p
                out[f'E_{k}_{j}_value'] = np.float64(J(p))
                out[f'E_{k}_{j}_grad'] = np.asarray(J.grad(p), dtype=float).reshape(-1)
                Hm = J.hessian(p)
                Hm = Hm.toarray() if hasattr(Hm, 'toarray') else np.asarray(Hm)
                # the reference returns [6x6 full | 0 ; m x 6 | lower-tri m x m] (dsm.py:369-371): keep its lower
                # triangle in full for small problems, and two symmetric mat-vec probes of it for all
                Hl = np.tril(Hm)
                Hs = Hl + np.tril(Hl, -1).T
                V = np.cos(np.outer(np.arange(1, 7 + m - 1 + 1), [0.7, 1.9]))
                out[f'E_{k}_{j}_hess_probe'] = Hs @ V
                if m <= 40 and j == 1: out[f'E_{k}_{j}_hess_lower'] = Hl
            # the reference's own driver on the restated solver
            with contextlib.redirect_stdout(io.StringIO()), contextlib.redirect_stderr(io.StringIO()):
                obj2 = ref.objects.Object()
                obj2.footprint = set(fp)
                cfg = {key: value for key, value in dsm_cfg.items() if key != 'smooth_mat_max_allocations'}
                cfg['cp_timeout'] = None
                res, fallback = ref.objects._compute_object(y_img, atoms, x_map, obj2, cfg, None)
            out[f'fit_{k}_energy'] = np.float64(res.energy)
            out[f'fit_{k}_on_boundary'] = np.bool_(res.on_boundary)
            out[f'fit_{k}_is_optimal'] = np.bool_(res.is_optimal)
            out[f'fit_{k}_fallback'] = np.bool_(fallback)
            out[f'fit_{k}_fg_offset'] = np.asarray(res.fg_offset, np.int64)
            out[f'fit_{k}_fg_fragment'] = np.asarray(res.fg_fragment, bool)
            print(name, k, sorted(fp), 'n', G.shape[0], 'm', m, 'energy', res.energy, 'opt', res.is_optimal, 'fallback', fallback)
        path = os.path.join(OUT, f'{name}.npz')
        np.savez_compressed(path, **out)
        print('wrote', path, os.path.getsize(path) // 1024, 'KiB')


if __name__ == '__main__':
    main()
