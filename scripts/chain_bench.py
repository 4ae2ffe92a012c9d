"""Link latency of single candidates: solves chosen candidates of the bench workload one at a time (nothing else on the GPU)."""
import sys, time
import numpy as np
from workloads import synth
from superdsm_b200.objects import get_engine

ids = [int(v) for v in sys.argv[1].split(',')] if len(sys.argv) > 1 else [202, 70, 31, 75, 550]
w = synth.make_workload('gowt1', oversegment=3, max_set_size=4, seed=1)
eng = get_engine(0)
eng.set_image(w['y'], w['y_mask'], w['atoms'], w['dsm_cfg']['background_margin'])
eng.set_params(w['dsm_cfg'])
for k in ids:
    fps = [w['footprints'][k]]
    eng.solve(fps)
    t0 = time.perf_counter(); res = eng.solve(fps); dt = time.perf_counter() - t0
    info = res.info[0]; st = eng.solver_stats(); ft = eng.finish_times_us()[0]
    ev = info[6] + info[7]
    print(f'cand {k} n {res.cands["n"][0]} m {info[0]} cells<= {res.cands["cell_cap"][0]} it {info[1:4]} nf {info[6]} nh {info[7]} solver {ft} us -> {ft / ev:.1f} us/link; '
          f'tasks eval {st["n_eval"]} hrows {st["n_hrows"]} ctrl {st["n_ctrl"]} cyc/ctrl-task {st["cyc_ctrl"] / max(st["n_ctrl"], 1):.0f} wall {dt * 1e3:.1f} ms')
