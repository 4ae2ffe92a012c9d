"""Debug aid: prints the device's and the oracle's interior-point iterates side by side for one golden candidate."""
import sys
import numpy as np
import scipy.sparse
sys.path.insert(0, '.')
from oracle import dsm_oracle as orc
from tests.golden_util import GoldenCase
from superdsm_b200.objects import get_engine

name, k = sys.argv[1], int(sys.argv[2])
case = GoldenCase(name)
eng = get_engine(0)
eng.set_image(case.y, case.y_mask, case.atoms, case.dsm_cfg['background_margin'])
eng.set_params(case.dsm_cfg)
res = eng.solve(case.footprints)
cfg = case.dsm_cfg
mask = case.mask(k); G = case.G(k); m = G.shape[1]
scale_cp = cfg['scale'] / mask.sum()
J_ell = orc.Energy(case.y, mask, cfg['epsilon'], cfg['alpha'], scipy.sparse.csr_matrix((int(mask.sum()), 0)))
J = orc.Energy(case.y, mask, cfg['epsilon'], cfg['alpha'], G)
tr = []
sol = orc._solve(J_ell, np.zeros(6), scale_cp, trace=tr)
theta = np.asarray(sol['x']).reshape(-1)
x0 = np.concatenate([theta, np.zeros(m)])
tr = []
sol = orc._solve(J, x0, scale_cp, trace=tr)
st, nit, x, dtr = eng.trace_solve(k, x0, use_xi=True)
ref = np.array([[t['f'], t['gap'], t['pres'], t['dres'], t['t'], t['s'], t['z'], t['relaxed']] for t in tr])
np.set_printoptions(linewidth=250, precision=12)
print('device status', st, 'iters', nit, 'oracle', sol['status'], sol['iterations'])
for i in range(max(len(ref), len(dtr))):
    a = dtr[i] if i < len(dtr) else None
    b = ref[i] if i < len(ref) else None
    print(i, 'dev', a)
    print(i, 'ora', b)
print('J dev', J(x), 'J ora', J(np.asarray(sol['x']).reshape(-1)))
