"""Debug aid: finish-time profile of the bench workload (per-link latency of the longest chains, busy CTAs over time)."""
import sys, time
import numpy as np
sys.path.insert(0, '.')
from workloads import synth
from superdsm_b200.objects import get_engine

w = synth.make_workload('gowt1', oversegment=3, max_set_size=4, seed=1)
eng = get_engine(0)
eng.set_image(w['y'], w['y_mask'], w['atoms'], w['dsm_cfg']['background_margin'])
eng.set_params(w['dsm_cfg'])
for _ in range(2):
    t0 = time.perf_counter(); res = eng.solve(w['footprints']); dt = time.perf_counter() - t0
print('solve', dt)
st = eng.solver_stats(); tot = max(st['cyc_total'], 1); wtot = max(st['cyc_wtotal'], 1)
print('ctrl CTAs: busy %.3f idle %.3f | eval warps: eval %.3f hrows %.3f idle %.3f finish %.3f' % (st['cyc_ctrl'] / tot, st['cyc_idle'] / tot, st['cyc_eval'] / wtot, st['cyc_hrows'] / wtot, st['cyc_widle'] / wtot, st['cyc_wfin'] / wtot),
      {k: st[k] for k in st if k.startswith('n_')}, 'urgent', st['q_tail_hi'], 'of', st['q_tail'])
print({k: (st['cyc_' + k] / max(st['n_' + k], 1)) for k in ('poly_f', 'poly_h', 'xi_f', 'xi_h')})
print('cycles per control task: big systems %.0f (%d), small %.0f (%d)' % (st['cyc_ctrl_big'] / max(st['n_ctrl_big'], 1), st['n_ctrl_big'], (st['cyc_ctrl'] - st['cyc_ctrl_big']) / max(st['n_ctrl'] - st['n_ctrl_big'], 1), st['n_ctrl'] - st['n_ctrl_big']))
print('big systems: cycles per factorisation %.0f, per (first) solve %.0f, factorisations %d' % (st['cyc_chol_big'] / max(st['n_chol_big'], 1), st['cyc_solve_big'] / max(st['n_chol_big'], 1), st['n_chol_big']))
info = res.info; ft = eng.finish_times_us().astype(float); ct = eng.control_times_us.astype(float); n = res.cands['n']; ev = (info[:, 6] + info[:, 7]).astype(float)
print('finish pct 10/25/50/75/90/99/100', np.percentile(ft, [10, 25, 50, 75, 90, 99, 100]))
hist, edges = np.histogram(ft, bins=20); print('finish histogram', hist.tolist(), edges[-1])
work = n * ev
o = np.argsort(ft)
print('cumulative pixel-evals share by finish time decile', [round(float(work[o[:int(len(o) * q / 10)]].sum() / work.sum()), 3) for q in range(1, 11)])
for k in np.argsort(-ft)[:15]:
    print('cand', k, 'n', n[k], 'm', info[k, 0], 'it', info[k, 1:4], 'nf', info[k, 6], 'nh', info[k, 7], 'done_us', ft[k], 'us/eval', round(ft[k] / max(ev[k], 1), 1), 'ctrl us/eval', round(ct[k] / max(ev[k], 1), 1))
for k in np.argsort(-n)[:8]:
    print('BIG cand', k, 'n', n[k], 'm', info[k, 0], 'it', info[k, 1:4], 'nf', info[k, 6], 'nh', info[k, 7], 'done_us', ft[k], 'us/eval', round(ft[k] / max(ev[k], 1), 1), 'ctrl us/eval', round(ct[k] / max(ev[k], 1), 1))
