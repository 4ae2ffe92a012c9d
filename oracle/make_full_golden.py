"""Generates ``tests/golden/full_<workload>.npz``: the oracle's fit of EVERY candidate of a BASELINE-shaped workload.

TEST INFRASTRUCTURE.  ``python -m oracle.make_full_golden gowt1 [processes]`` (minutes to an hour of CPU per workload;
the workloads are the ones ``bench.py --workload`` times, see ``WORKLOAD_ARGS``).  Per candidate k the fixture holds what
``tests/test_gpu_full_parity.py`` compares the CUDA path with:

* ``energy``, ``status`` (0 optimal, 1 fallback, 2 skipped: one foreground pixel, 3 CvxprogError), ``on_boundary``,
* ``mask_hash``: first 8 bytes of the SHA-1 of (fg_offset, fg_fragment shape, packed fg_fragment) -- objects.py:198-211,
* ``params`` (theta, xi; ragged, ``par_off``), ``n``, ``m``,
* ``iters`` (elliptical pass 1, pass 2, DSM solve; -1: not run), ``cp_status`` of the DSM solve (0 optimal, 1 unknown, 2 raised),
* ``max_relaxed``: the largest relaxed-iteration counter the DSM solve went through, ``chol_fails``: failed factorisations in it
  (a "regular" path has max_relaxed <= 1 and no failed factorisation: only those are compared parameter by parameter).
"""
import hashlib
import os
import sys
import time

for _v in ('OMP_NUM_THREADS', 'OPENBLAS_NUM_THREADS', 'MKL_NUM_THREADS'): os.environ[_v] = '1'

import numpy as np                                  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path: sys.path.insert(0, ROOT)

OUT = os.path.join(ROOT, 'tests', 'golden')

#: workload name -> arguments of tests.workloads.make_workload (the same lists bench.py --workload times)
WORKLOAD_ARGS = {
    'gowt1':   dict(name='gowt1', oversegment=3, max_set_size=4, seed=1),
    'bbbc039': dict(name='bbbc039', oversegment=3, max_set_size=4, seed=39),
    'nih3t3':  dict(name='nih3t3', oversegment=3, max_set_size=4, seed=3),
    'tiled':   dict(name='tiled4k', oversegment=3, max_set_size=4, seed=5, shape=(1024, 1024), n=1250, max_candidates=8000),
}


def mask_hash(fg_offset, fg_fragment):
    h = hashlib.sha1()
    h.update(np.asarray(fg_offset, np.int64).tobytes())
    h.update(np.asarray(fg_fragment.shape, np.int64).tobytes())
    h.update(np.packbits(np.asarray(fg_fragment, bool)).tobytes())
    return np.frombuffer(h.digest()[:8], dtype=np.uint64)[0]


_w = None


def _init(wname):
    global _w
    from workloads.synth import make_workload
    _w = make_workload(**WORKLOAD_ARGS[wname])


def _fit(k):
    from oracle import dsm_oracle as orc
    info = {}
    t0 = time.perf_counter()
    try:
        r = orc.compute_object(_w['y'], _w['y_mask'], _w['atoms'], _w['footprints'][k], _w['dsm_cfg'], info=info)
    except orc.CvxprogError:
        return dict(k=k, status=3, energy=0.0, seconds=time.perf_counter() - t0)
    if r['params'] is None:
        return dict(k=k, status=2, energy=0.0, seconds=time.perf_counter() - t0, n=int(r['mask'].sum()))
    it1 = info['pass1'][1] if info.get('pass1') else -1
    it2 = info['pass2'][1] if info.get('pass2') else -1
    if 'dsm' in info: cp_status, it3 = (0 if info['dsm'][0] == 'optimal' else 1), info['dsm'][1]
    else: cp_status, it3 = 2, len(info.get('dsm_trace') or []) - 1
    tr = info.get('dsm_trace') or []
    return dict(k=k, status=1 if r['fallback'] else 0, energy=float(r['energy']), on_boundary=bool(r['on_boundary']),
                mask_hash=mask_hash(r['fg_offset'], r['fg_fragment']), params=np.asarray(r['params'], float), n=int(r['n']), m=int(r['m']),
                iters=(it1, it2, it3), cp_status=cp_status, max_relaxed=max([abs(t['relaxed']) for t in tr], default=0),
                chol_fails=len(info.get('dsm_events') or []), error=info.get('dsm_error', ''), seconds=time.perf_counter() - t0)


def main():
    import multiprocessing as mp
    wname = sys.argv[1]
    procs = int(sys.argv[2]) if len(sys.argv) > 2 else (os.cpu_count() or 1)
    _init(wname)
    K = len(_w['footprints'])
    t0 = time.time()
    with mp.get_context('fork').Pool(procs, initializer=_init, initargs=(wname,)) as pool:
        res = pool.map(_fit, range(K), chunksize=2)
    secs = time.time() - t0
    energy = np.zeros(K); status = np.zeros(K, np.int8); onb = np.zeros(K, bool); mh = np.zeros(K, np.uint64)
    n = np.zeros(K, np.int32); m = np.zeros(K, np.int32); iters = -np.ones((K, 3), np.int16); cp_status = -np.ones(K, np.int8)
    max_relaxed = np.zeros(K, np.int8); chol_fails = np.zeros(K, np.int16); cpu_s = np.zeros(K, np.float32)
    par_off = np.zeros(K + 1, np.int64); params = []
    errors = []
    for r in res:
        k = r['k']
        status[k], energy[k], cpu_s[k], n[k] = r['status'], r['energy'], r['seconds'], r.get('n', 0)
        par_off[k + 1] = par_off[k]
        if r['status'] >= 2: continue
        onb[k], mh[k], m[k], iters[k], cp_status[k] = r['on_boundary'], r['mask_hash'], r['m'], r['iters'], r['cp_status']
        max_relaxed[k], chol_fails[k] = r['max_relaxed'], r['chol_fails']
        params.append(r['params'])
        par_off[k + 1] = par_off[k] + len(r['params'])
        if r['error']: errors.append(f"{k}: {r['error']}")
    fp_len = np.array([len(fp) for fp in _w['footprints']], np.int32)
    fp_flat = np.concatenate([sorted(fp) for fp in _w['footprints']]).astype(np.int32)
    os.makedirs(OUT, exist_ok=True)
    path = os.path.join(OUT, f'full_{wname}.npz')
    np.savez_compressed(path, energy=energy, status=status, on_boundary=onb, mask_hash=mh, n=n, m=m, iters=iters, cp_status=cp_status,
                        max_relaxed=max_relaxed, chol_fails=chol_fails, cpu_seconds=cpu_s, par_off=par_off,
                        params=np.concatenate(params) if params else np.zeros(0), fp_len=fp_len, fp_flat=fp_flat,
                        errors=np.array(errors), image_checksum=np.float64(_w['y'].sum()), atoms_checksum=np.int64(_w['atoms'].astype(np.int64).sum()),
                        restore_policy=np.array(__import__('oracle.cpsolver', fromlist=['x']).RESTORE_POLICY))
    print(f'wrote {path}: {K} candidates, {secs:.0f} s wall on {procs} processes, {cpu_s.sum():.0f} core-s, '
          f'status counts {np.bincount(status, minlength=4).tolist()}, irregular {(int((max_relaxed > 1).sum()))}, '
          f'chol failures in {(int((chol_fails > 0).sum()))}, {os.path.getsize(path) // 1024} KiB')
    for e in errors[:20]: print('  dsm solve raised:', e)


if __name__ == '__main__':
    main()
