"""Line-search statistics of the oracle on candidates of a workload (CPU): how many trials per pass, and how often the trial
points of the second pass repeat those of the first (same first step, same direction)."""
import os, sys
for _v in ('OMP_NUM_THREADS', 'OPENBLAS_NUM_THREADS', 'MKL_NUM_THREADS'): os.environ[_v] = '1'
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import cpsolver, dsm_oracle as orc
from oracle.make_full_golden import WORKLOAD_ARGS
from workloads.synth import make_workload

w = make_workload(**WORKLOAD_ARGS[sys.argv[1] if len(sys.argv) > 1 else 'gowt1'])
z = np.load(f'tests/golden/full_{sys.argv[1] if len(sys.argv) > 1 else "gowt1"}.npz')
order = np.argsort(-z['cpu_seconds'])
picks = list(order[:6]) + list(order[len(order) // 4::len(order) // 12][:8])
tot = dict(ls=0, trials=0, p0=0, p1=0, same_first=0, p1_hits=0, p1_trials=0, links_now=0, links_cache=0, links_spec=0, evals_spec=0)
for k in picks:
    cpsolver.LS_LOG = []
    try: orc.compute_object(w['y'], w['y_mask'], w['atoms'], w['footprints'][k], w['dsm_cfg'])
    except orc.CvxprogError: continue
    log = cpsolver.LS_LOG
    # solves are concatenated; pair pass 0 / pass 1 entries
    i = 0
    last = {0: 1, 1: 1}
    c = dict(tot); [c.__setitem__(kk, 0) for kk in c]
    while i + 1 < len(log):
        a, b = log[i], log[i + 1]
        if a[1] != 0 or b[1] != 1: i += 1; continue
        i += 2
        c['ls'] += 2; c['p0'] += a[3]; c['p1'] += b[3]; c['trials'] += a[3] + b[3]
        same = a[2] == b[2]
        c['same_first'] += same
        hits = min(a[3], b[3]) if same else 0
        c['p1_hits'] += hits; c['p1_trials'] += b[3]
        c['links_now'] += a[3] + b[3] + 1
        c['links_cache'] += a[3] + (b[3] - hits) + 1
        # speculation: first request K = trials of the previous search of the same pass (<= 8), then doubling
        def links(T, K0, have):
            n, got, K, ev = 0, have, max(1, min(8, K0)), 0
            while got < T:
                need = K if got + K <= 8 or got >= 8 else max(1, 8 - got)
                n += 1; got += need; ev += need; K = min(8, 2 * K)
            return n, ev
        l0, e0 = links(a[3], last[0], 0)
        stored = max(a[3], e0) if same else 0
        l1, e1 = links(b[3], last[1], stored)
        last[0], last[1] = a[3], b[3]
        c['links_spec'] += l0 + l1 + 1; c['evals_spec'] += e0 + e1
    print(f"cand {k} n {z['n'][k]} m {z['m'][k]} it {z['iters'][k].tolist()}: searches {c['ls']} trials/search p0 {c['p0'] / max(1, c['ls'] // 2):.2f} p1 {c['p1'] / max(1, c['ls'] // 2):.2f} "
          f"same first step {c['same_first']}/{c['ls'] // 2} p1 hits {c['p1_hits']}/{c['p1_trials']} links now {c['links_now']} cache {c['links_cache']} spec {c['links_spec']} "
          f"evals now {c['trials']} spec {c['evals_spec']}", flush=True)
    for kk in tot: tot[kk] += c[kk]
print('total', tot)
