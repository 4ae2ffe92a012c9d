"""How far is the solve from the pure evaluation throughput?  pixel-evaluations of one solve / sweep rate (scripts/sweep_bench.py)."""
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
print('info columns sample', info[:3])
px = n.sum()
f_ms = eng.sweep_eval_ms(h, repeats=5)
h_ms = eng.sweep_eval_ms(h, repeats=5, want_h=True)
r_ms = eng.sweep_eval_ms(h, repeats=5, want_h=True, rows=True)
nf = info[:, 6].astype(np.float64); nh = info[:, 7].astype(np.float64)
print('sweeps ms', f_ms, h_ms, r_ms, 'pixels', px)
print('nf sum', nf.sum(), 'nh sum', nh.sum())
t = ((n * nf).sum() * f_ms + (n * nh).sum() * (h_ms + r_ms)) / px
print('pixel-evals f', (n * nf).sum(), 'h', (n * nh).sum(), 'ideal ms at sweep rate', t)
