"""Every candidate of a bench workload, device result against the oracle's fit (TEST INFRASTRUCTURE: imports oracle/).

    python scripts/full_oracle_compare.py [workload] [processes]  ->  one summary line + the candidates that differ
"""
import sys, os, json, time
sys.path.insert(0, '.')
import numpy as np
import multiprocessing as mp

name = sys.argv[1] if len(sys.argv) > 1 else 'gowt1'
procs = int(sys.argv[2]) if len(sys.argv) > 2 else min(32, os.cpu_count() or 1)
_w = None


def _init():
    global _w
    from workloads import synth
    os.environ['OMP_NUM_THREADS'] = '1'
    _w = synth.make_workload(name, oversegment=3, max_set_size=4, seed=1)


def _fit(k):
    from oracle import dsm_oracle as orc
    try:
        r = orc.compute_object(_w['y'], _w['y_mask'], _w['atoms'], _w['footprints'][k], _w['dsm_cfg'])
        return k, r['energy'], r['is_optimal'], r['fallback'], r['fg_offset'], r['fg_fragment'], r.get('n', 0), None
    except orc.CvxprogError as e:
        return k, np.nan, False, False, None, None, 0, 'CvxprogError'


if __name__ == '__main__':
    from superdsm_b200 import _lib
    from workloads import synth
    from superdsm_b200.objects import get_engine
    w = synth.make_workload(name, oversegment=3, max_set_size=4, seed=1)
    eng = get_engine(0)
    eng.set_image(w['y'], w['y_mask'], w['atoms'], w['dsm_cfg']['background_margin'])
    eng.set_params(w['dsm_cfg'])
    res = eng.solve(w['footprints'])
    t0 = time.time()
    with mp.get_context('fork').Pool(procs, initializer=_init) as pool:
        out = pool.map(_fit, range(len(w['footprints'])), chunksize=4)
    secs = time.time() - t0
    n_ok = n_mask = n_status = 0
    bad = []
    for k, e, opt, fb, off, frag, n, err in out:
        st = int(res.status[k])
        if err is not None:
            same = st == _lib.ST_ERROR
            n_ok += same; n_mask += same; n_status += same
            if not same: bad.append((k, 'oracle error', st))
            continue
        tol = 1e-6 * abs(e) + 2e-7 * max(n, 1) / w['dsm_cfg']['scale']
        e_ok = abs(res.energy[k] - e) <= tol
        d_off, d_frag = res.foreground(int(k))
        m_ok = np.array_equal(d_frag, frag) and np.array_equal(d_off, off)
        s_ok = (st == _lib.ST_OPTIMAL) == bool(opt) and (st == _lib.ST_FALLBACK) == bool(fb)
        n_ok += e_ok; n_mask += m_ok; n_status += s_ok
        if not (e_ok and m_ok and s_ok): bad.append((int(k), float(res.energy[k]), float(e), st, bool(opt), bool(fb), bool(m_ok)))
    total = len(out)
    print(json.dumps(dict(workload=name, candidates=total, energy_within_tolerance=int(n_ok), foreground_identical=int(n_mask),
                          status_equal=int(n_status), oracle_seconds=round(secs, 1), processes=procs,
                          tolerance='1e-6 |E| + 2e-7 n / scale')))
    for b in bad[:40]: print(' differs:', b)
