"""Access to the committed golden fixtures (made by oracle/make_golden.py from the reference's own code)."""
import os

import numpy as np
import scipy.sparse

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
CASES = ('toy', 'odd', 'hand')


class GoldenCase(object):
    def __init__(self, name):
        self.name = name
        self.z = np.load(os.path.join(GOLDEN_DIR, name + '.npz'))
        self.y = self.z['y']
        self.y_mask = self.z['y_mask']
        self.atoms = self.z['atoms']
        lens = self.z['fp_len']
        flat = self.z['fp_flat']
        offs = np.concatenate([[0], np.cumsum(lens)])
        self.footprints = [frozenset(int(a) for a in flat[offs[k]:offs[k + 1]]) for k in range(len(lens))]
        self.dsm_cfg = {key[4:]: float(self.z[key]) for key in self.z.files if key.startswith('cfg_')}
        for key in ('smooth_subsample', 'background_margin', 'cachesize'):
            self.dsm_cfg[key] = int(self.dsm_cfg[key])
        if self.dsm_cfg['smooth_amount'] == int(self.dsm_cfg['smooth_amount']):
            self.dsm_cfg['smooth_amount'] = int(self.dsm_cfg['smooth_amount'])
        self.dsm_cfg.update(init='elliptical', smooth_mat_dtype='float32', cachetest=None)

    def __len__(self): return len(self.footprints)

    def mask(self, k):
        return np.unpackbits(self.z[f'mask_{k}'])[:self.y.size].reshape(self.y.shape).astype(bool)

    def G(self, k):
        shape = tuple(int(v) for v in self.z[f'G_{k}_shape'])
        return scipy.sparse.csr_matrix((self.z[f'G_{k}_data'].astype(np.float64), self.z[f'G_{k}_indices'], self.z[f'G_{k}_indptr']), shape=shape)

    def energy_points(self, k):
        for j in range(3):
            yield dict(p=self.z[f'E_{k}_{j}_p'], value=float(self.z[f'E_{k}_{j}_value']), grad=self.z[f'E_{k}_{j}_grad'],
                       hess_probe=self.z[f'E_{k}_{j}_hess_probe'],
                       hess_lower=self.z[f'E_{k}_{j}_hess_lower'] if f'E_{k}_{j}_hess_lower' in self.z.files else None)

    def fit(self, k):
        return {key: self.z[f'fit_{k}_{key}'] for key in ('energy', 'on_boundary', 'is_optimal', 'fallback', 'fg_offset', 'fg_fragment')}


def probe_matrix(N):
    return np.cos(np.outer(np.arange(1, N + 1), [0.7, 1.9]))
