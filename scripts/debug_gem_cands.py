import sys
sys.path.insert(0, '.')
import numpy as np
from tests.test_gem_seam import _load
from oracle.make_gem_golden import GEM_ARGS
from superdsm_b200.objects import get_engine
from workloads.synth import make_workload
name = 'gowt1'
z, fps, cover = _load(name)
w = make_workload(**GEM_ARGS[name])
eng = get_engine(0)
eng.set_image(w['y'], w['y_mask'], w['atoms'], w['dsm_cfg']['background_margin'])
eng.set_params(w['dsm_cfg'])
ks = [int(a) for a in sys.argv[1:]] or [670, 693, 1025, 1083, 1780]
res = eng.solve([fps[k] for k in ks])
for i, k in enumerate(ks):
    print('k', k, 'n', int(res.cands['n'][i]), 'status', int(res.status[i]), 'E', float(res.energy[i]), 'info [m it1 it2 it3 nfg bits nf nh]', res.info[i].tolist(), 'oracle E', float(z['energy'][k]))
