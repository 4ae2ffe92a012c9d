"""Comparison of the CUDA path with the committed full-config oracle fixtures (``tests/golden/full_<workload>.npz``,
made by ``oracle/make_full_golden.py``): every candidate of a BASELINE-shaped workload.  TEST INFRASTRUCTURE."""
import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')


def load_fixture(wname):
    return np.load(os.path.join(GOLDEN_DIR, f'full_{wname}.npz'))


def make_workload(wname):
    from oracle.make_full_golden import WORKLOAD_ARGS
    from workloads.synth import make_workload as mk
    return mk(**WORKLOAD_ARGS[wname])


def compare(eng, wname, max_candidates=None):
    """Solves every candidate of the workload on the device and compares with the fixture.  Returns a dict of per-candidate
    arrays: ``status_equal``, ``mask_equal``, ``energy_err`` (|E_dev - E_ora| minus the absolute floor 2e-7 n / scale, relative to
    |E_ora|), ``param_err`` (max |p_dev - p_ora| / max(1, max |p_ora|); nan where not comparable), ``regular``, ..."""
    from oracle.make_full_golden import mask_hash
    from superdsm_b200 import _lib
    z = load_fixture(wname)
    w = make_workload(wname)
    assert float(w['y'].sum()) == float(z['image_checksum']) and int(w['atoms'].astype(np.int64).sum()) == int(z['atoms_checksum']), \
        'the synthetic workload is not the one the fixture was made from'
    fps = w['footprints']
    K = len(fps) if max_candidates is None else min(len(fps), max_candidates)
    assert len(fps) == len(z['energy'])
    eng.set_image(w['y'], w['y_mask'], w['atoms'], w['dsm_cfg']['background_margin'])
    eng.set_params(w['dsm_cfg'])
    res = eng.solve(fps[:K])
    scale = float(w['dsm_cfg']['scale'])
    o_status, o_energy, o_hash = z['status'][:K], z['energy'][:K], z['mask_hash'][:K]
    d_status = res.status.astype(np.int64)
    status_equal = d_status == o_status
    mask_equal = np.zeros(K, bool)
    param_err = np.full(K, np.nan)
    energy_err = np.zeros(K)
    n = res.cands['n'].astype(float)
    for k in range(K):
        if d_status[k] >= _lib.ST_SKIPPED or o_status[k] >= 2:
            mask_equal[k] = status_equal[k]
            continue
        off, frag = res.foreground(k)
        mask_equal[k] = mask_hash(off, frag) == o_hash[k]
        floor = 2e-7 * n[k] / scale                  # cvxopt's abstol on the scaled objective scale / n * psi
        energy_err[k] = max(abs(res.energy[k] - o_energy[k]) - floor, 0.0) / max(abs(o_energy[k]), 1e-300)
        p_o = z['params'][z['par_off'][k]:z['par_off'][k + 1]]
        p_d = res.params(k)
        if len(p_o) == len(p_d): param_err[k] = np.abs(p_d - p_o).max() / max(1.0, np.abs(p_o).max())
    regular = (z['max_relaxed'][:K] <= 1) & (z['chol_fails'][:K] == 0) & (z['cp_status'][:K] == 0)
    return dict(K=K, res=res, fixture=z, status_equal=status_equal, mask_equal=mask_equal, energy_err=energy_err, param_err=param_err,
                regular=regular, d_status=d_status, o_status=o_status, d_energy=res.energy[:K].copy(), o_energy=o_energy, n=n,
                d_iters=res.info[:K, 1:4].copy(), o_iters=z['iters'][:K], m_equal=res.info[:K, 0] == np.where(o_status < 2, z['m'][:K], res.info[:K, 0]))


def summary(c):
    K = c['K']
    both_opt = (c['d_status'] == 0) & (c['o_status'] == 0)
    reg = c['regular'] & both_opt
    pe = c['param_err']
    out = dict(candidates=K, status_equal=int(c['status_equal'].sum()), mask_equal=int(c['mask_equal'].sum()),
               energy_within_1e6=int((c['energy_err'] <= 1e-6).sum()), energy_worst=float(c['energy_err'].max()),
               energy_gross_gt_1e3=np.nonzero(c['energy_err'] > 1e-3)[0].tolist(),
               device_above_oracle=np.nonzero((c['d_energy'] - c['o_energy']) > 1e-6 * np.abs(c['o_energy']) + 2e-7 * c['n'] / 1000.0)[0].tolist()[:50],
               regular=int(reg.sum()), regular_param_worst=float(np.nanmax(pe[reg])) if reg.any() else 0.0,
               regular_param_gt_1e6=int((pe[reg] > 1e-6).sum()), all_param_median=float(np.nanmedian(pe[both_opt])) if both_opt.any() else 0.0,
               all_param_gt_1e3=int((pe[both_opt] > 1e-3).sum()), m_equal=int(c['m_equal'].sum()))
    return out
