"""Debug aid: device and oracle interior-point iterates side by side for one candidate of the bench workload."""
import sys
import numpy as np
import scipy.sparse
sys.path.insert(0, '.')
from oracle import dsm_oracle as orc
from workloads import synth
from superdsm_b200.objects import get_engine

k = int(sys.argv[1])
w = synth.make_workload('gowt1', oversegment=3, max_set_size=4, seed=1)
cfg = dict(w['dsm_cfg'])
eng = get_engine(0)
eng.set_image(w['y'], w['y_mask'], w['atoms'], cfg['background_margin'])
eng.set_params(cfg)
res = eng.solve(w['footprints'])
print('batch result: status', res.status[k], 'energy', res.energy[k], 'info', res.info[k])
mask = orc.cvxprog_region_mask(w['y'], w['y_mask'], w['atoms'], w['footprints'][k], cfg['background_margin'])
G = orc.smooth_matrix(mask, cfg['smooth_amount'], cfg['gaussian_shape_multiplier'], cfg['smooth_subsample'])
m = G.shape[1]
scale_cp = cfg['scale'] / mask.sum()
J_ell = orc.Energy(w['y'], mask, cfg['epsilon'], cfg['alpha'], scipy.sparse.csr_matrix((int(mask.sum()), 0)))
J = orc.Energy(w['y'], mask, cfg['epsilon'], cfg['alpha'], G)
theta = orc.elliptical_solution(J_ell, scale_cp)
x0 = np.concatenate([theta, np.zeros(m)])
tr = []
sol = orc._solve(J, x0, scale_cp, trace=tr)
st, nit, x, dtr = eng.trace_solve(k, x0, use_xi=True)
ref = np.array([[t['f'], t['gap'], t['pres'], t['dres'], t['t'], t['s'], t['z'], t['relaxed']] for t in tr])
np.set_printoptions(linewidth=250, precision=14)
print('device status', st, 'iters', nit, 'oracle', sol['status'], sol['iterations'])
for i in range(max(len(ref), len(dtr))):
    print(i, 'dev', dtr[i] if i < len(dtr) else None)
    print(i, 'ora', ref[i] if i < len(ref) else None)
print('J dev', J(x), 'J ora', J(np.asarray(sol['x']).reshape(-1)))
