"""Device vs the committed full-config oracle fixtures, every candidate: python scripts/full_parity_report.py [workload ...]"""
import json, sys
sys.path.insert(0, '.')
import numpy as np
from tests import full_parity_util as fpu
from superdsm_b200.objects import get_engine

eng = get_engine(0)
for wname in (sys.argv[1:] or ['gowt1', 'bbbc039', 'nih3t3', 'tiled']):
    try:
        c = fpu.compare(eng, wname)
    except FileNotFoundError as e:
        print(wname, 'no fixture', e); continue
    s = fpu.summary(c)
    np.savez_compressed(f'gpurun_out/parity_{wname}.npz', **{k: v for k, v in c.items() if isinstance(v, np.ndarray)}, d_info=c['res'].info)
    print(wname, json.dumps(s))
    bad = np.nonzero(~c['status_equal'] | ~c['mask_equal'] | (c['energy_err'] > 1e-6))[0]
    for k in bad[:60]:
        print(f"  {wname} k={k} n={int(c['n'][k])} status d/o {c['d_status'][k]}/{c['o_status'][k]} mask_eq {bool(c['mask_equal'][k])} E d/o {c['d_energy'][k]:.9g}/{c['o_energy'][k]:.9g} "
              f"err {c['energy_err'][k]:.2e} iters d {c['d_iters'][k].tolist()} o {c['o_iters'][k].tolist()} cp_status {int(c['fixture']['cp_status'][k])} "
              f"max_relaxed {int(c['fixture']['max_relaxed'][k])} chol_fails {int(c['fixture']['chol_fails'][k])} bits {int(c['res'].info[k, 5])}")
