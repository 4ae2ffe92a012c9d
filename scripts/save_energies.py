"""Debug aid: solves the bench workload and stores / compares the per-candidate results (A/B check of solver changes)."""
import sys
import numpy as np
sys.path.insert(0, '.')
from workloads import synth
from superdsm_b200.objects import get_engine
path = sys.argv[1]
w = synth.make_workload('gowt1', oversegment=3, max_set_size=4, seed=1)
eng = get_engine(0)
eng.set_image(w['y'], w['y_mask'], w['atoms'], w['dsm_cfg']['background_margin'])
eng.set_params(w['dsm_cfg'])
res = eng.solve(w['footprints'])
if len(sys.argv) > 2 and sys.argv[2] == 'compare':
    ref = np.load(path)
    e0, e1 = ref['energy'], res.energy
    rel = np.abs(e1 - e0) / np.maximum(np.abs(e0), 1e-300)
    n = res.cands['n']
    tol = 1e-6 * np.abs(e0) + 2e-7 * n / w['dsm_cfg']['scale']
    print('status equal', np.array_equal(ref['status'], res.status), 'max rel energy diff', rel.max(), 'outside tolerance', int((np.abs(e1 - e0) > tol).sum()),
          'iterations equal', float((ref['info'][:, 1:4] == res.info[:, 1:4]).all(axis=1).mean()))
    bad = np.where(np.abs(e1 - e0) > tol)[0][:10]
    for k in bad: print(' cand', k, 'n', n[k], 'm', res.info[k, 0], e0[k], e1[k], ref['info'][k, 1:4], res.info[k, 1:4])
else:
    np.savez(path, energy=res.energy, status=res.status, info=res.info)
    print('saved', path, len(res.energy))
