"""Times the evaluation sweeps (value+gradient, +Hessian, Hessian rows) over the whole bench workload, outside the solver."""
import sys
sys.path.insert(0, ".")
import numpy as np
from workloads import synth
from superdsm_b200.objects import get_engine

name = sys.argv[1] if len(sys.argv) > 1 else 'gowt1'
w = synth.make_workload(name, oversegment=3, max_set_size=4, seed=1)
eng = get_engine(0)
eng.set_image(w['y'], w['y_mask'], w['atoms'], w['dsm_cfg']['background_margin'])
eng.set_params(w['dsm_cfg'])
h = eng.prepare(w['footprints'])
eng.run_prepared(h)
eng.torch.cuda.synchronize()
n = h['cands']['n'].astype(np.float64)
info = h['out']['info'][:32 * h['n']].cpu().numpy().view(np.int32).reshape(-1, 8)
m = info[:, 0]
px = n.sum()
for label, kw in (('value+grad', dict()), ('value+grad+cell Hessians', dict(want_h=True)), ('Hessian rows', dict(want_h=True, rows=True))):
    ms = eng.sweep_eval_ms(h, repeats=5, **kw)
    print(f'{label:28s} {ms:8.3f} ms  {px / ms * 1e-6:8.2f} Gpx/s  alg {12 * px / ms * 1e-6:8.1f} GB/s  clk/px/SM {ms * 1e-3 * 1.965e9 * 148 / px:6.2f}')
print('pixels', int(px), 'candidates', h['n'], 'sum m', int(m.sum()))
