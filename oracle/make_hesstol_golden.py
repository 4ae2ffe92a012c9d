"""``tests/golden/hesstol.npz``: the reference's ``Energy.hessian`` with ``hessian_sparsity_tol > 0`` (dsm.py:376-382) on candidates of
the committed golden cases.  TEST INFRASTRUCTURE, build container only: ``python -m oracle.make_hesstol_golden``."""
import contextlib
import io
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_harness                     # noqa: E402
from tests.golden_util import GoldenCase           # noqa: E402


def main():
    ref = ref_harness.load()
    out = {}
    for name, ks in (('hand', (0, 2)), ('toy', (0, 3))):
        case = GoldenCase(name)
        cfg = case.dsm_cfg
        y_img = ref.image.Image.create_from_array(case.y, normalize=False, mask=case.y_mask)
        for k in ks:
            obj = ref.objects.Object()
            obj.footprint = set(case.footprints[k])
            region = obj.get_cvxprog_region(y_img, case.atoms, cfg['background_margin'])
            p = list(case.energy_points(k))[1]['p']
            mk = lambda tol: ref.dsm.Energy(region, cfg['epsilon'], cfg['alpha'], ref.dsm.SmoothMatrixFactory(
                cfg['smooth_amount'], cfg['gaussian_shape_multiplier'], cfg['smooth_subsample'], None, 'float32'), 0, tol)
            with contextlib.redirect_stdout(io.StringIO()):
                H0 = mk(0).hessian(p)
                H0 = np.tril(H0.toarray() if hasattr(H0, 'toarray') else np.asarray(H0))
                off = np.abs(H0[np.tril_indices_from(H0, -1)])
                tol = float(np.median(off[off > 0])) * 1.001          # not exactly an entry's value: the comparison must not hinge on the last bit
                H1 = mk(tol).hessian(p)
                H1 = np.tril(H1.toarray() if hasattr(H1, 'toarray') else np.asarray(H1))
            out[f'{name}_{k}_tol'] = np.float64(tol)
            out[f'{name}_{k}_p'] = p
            out[f'{name}_{k}_hess_lower'] = H1
            print(name, k, 'N', H1.shape[0], 'tol', tol, 'non-zeros', int((H0 != 0).sum()), '->', int((H1 != 0).sum()))
    path = os.path.join(ROOT, 'tests', 'golden', 'hesstol.npz')
    np.savez_compressed(path, **out)
    print('wrote', path, os.path.getsize(path) // 1024, 'KiB')


if __name__ == '__main__':
    main()
