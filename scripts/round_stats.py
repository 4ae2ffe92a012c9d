"""Debug aid: solves the bench workload once and stores the per-candidate counters (iterations, evaluations, sizes)."""
import sys, time
import numpy as np
sys.path.insert(0, '.')
from workloads import synth
from superdsm_b200.objects import get_engine

name = sys.argv[1] if len(sys.argv) > 1 else 'gowt1'
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
w = synth.make_workload(name, oversegment=3, max_set_size=4, seed=1)
eng = get_engine(0)
eng.set_image(w['y'], w['y_mask'], w['atoms'], w['dsm_cfg']['background_margin'])
eng.set_params(w['dsm_cfg'])
for _ in range(reps):
    t0 = time.perf_counter()
    res = eng.solve(w['footprints'])
    dt = time.perf_counter() - t0
print('solve', dt, 's launches', eng.launch_count)
st = eng.solver_stats()
print('solver stats', st)
tot = max(st['cyc_total'], 1)
print('CTA time share: eval %.3f hrows %.3f ctrl %.3f idle %.3f' % tuple(st[k] / tot for k in ('cyc_eval', 'cyc_hrows', 'cyc_ctrl', 'cyc_idle')))
info = res.info
ft = eng.finish_times_us()
print('finish time (us) percentiles 50/90/99/100', np.percentile(ft, [50, 90, 99, 100]))
n = res.cands['n']
ev = info[:, 6] + info[:, 7]
print('status counts', np.bincount(res.status))
print('evals per candidate: percentiles 0/50/90/99/100', np.percentile(ev, [0, 50, 90, 99, 100]))
print('n percentiles', np.percentile(n, [0, 50, 90, 99, 100]), 'm percentiles', np.percentile(info[:, 0], [0, 50, 90, 99, 100]))
order = np.argsort(-ft)[:25]
for k in order:
    print('cand', k, 'n', n[k], 'm', info[k, 0], 'it', info[k, 1:4], 'n_fg', info[k, 4], 'bits', info[k, 5], 'nf', info[k, 6], 'nh', info[k, 7], 'status', res.status[k], 'E', res.energy[k], 'done_us', ft[k])
np.savez('gpurun_out/round_stats.npz', info=info, n=n, status=res.status, energy=res.energy, cell_cap=res.cands['cell_cap'])
