"""Candidate throughput on a crop of the 4096 x 4096 tiled field (BASELINE config 4): many small candidates."""
import sys, time
import numpy as np
sys.path.insert(0, '.')
from superdsm_b200 import _lib
    from workloads import synth
from superdsm_b200.objects import get_engine
side = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
n = int(20000 * (side / 4096) ** 2)
t0 = time.perf_counter()
w = synth.make_workload('tiled4k', oversegment=2, max_set_size=3, shape=(side, side), n=n)
print('workload', side, n, 'nuclei', len(w['footprints']), 'candidates', round(time.perf_counter() - t0, 1), 's')
eng = get_engine(0)
eng.set_image(w['y'], w['y_mask'], w['atoms'], w['dsm_cfg']['background_margin'])
eng.set_params(w['dsm_cfg'])
h = eng.prepare(w['footprints'])
import torch
for rep in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter(); eng.run_prepared(h); torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print('solve %.3f s  %.0f candidate-solves/s' % (dt, h['n'] / dt))
st = eng.solver_stats(h['ws'])
status = h['out']['status'][:4 * h['n']].cpu().numpy().view(np.int32)
print('done', st['done'], 'abort', st['abort'], 'status counts', np.bincount(status), 'pixels', int(h['cands']['n'].sum()), 'workspace GB', h['ws'].numel() / 1e9)
wt = max(st['cyc_wtotal'], 1)
print('eval warps: eval %.3f hrows %.3f idle %.3f | ctrl CTAs busy %.3f' % (st['cyc_eval'] / wt, st['cyc_hrows'] / wt, st['cyc_widle'] / wt, st['cyc_ctrl'] / max(st['cyc_total'], 1)))
