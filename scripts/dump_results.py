"""Solves whole workloads with the library SDSM_LIB names (and the SDSM_* tuning variables of the environment) and dumps every
result array: ``python scripts/dump_results.py out.npz [workload ...]``; ``--compare a.npz b.npz`` checks two dumps for bitwise
equality (energies, parameters, status, iteration / evaluation counters, foreground masks)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np


def dump(path, names):
    import torch
    from oracle.make_full_golden import WORKLOAD_ARGS
    from workloads.synth import make_workload
    from superdsm_b200.engine import Engine
    eng = Engine(0)
    out = {}
    for name in names:
        w = make_workload(**WORKLOAD_ARGS[name])
        eng.set_image(w['y'], w['y_mask'], w['atoms'], w['dsm_cfg']['background_margin'])
        eng.set_params(w['dsm_cfg'])
        res = eng.solve(w['footprints'])
        ts = []
        for _ in range(3):
            t0 = time.perf_counter(); res = eng.solve(w['footprints']); ts.append(time.perf_counter() - t0)
        K = len(w['footprints'])
        out[f'{name}_energy'] = res.energy.copy(); out[f'{name}_status'] = res.status.copy(); out[f'{name}_info'] = res.info.copy()
        out[f'{name}_params'] = np.concatenate([res.params(k) if res.status[k] < 2 else np.zeros(0) for k in range(K)])
        out[f'{name}_fgbox'] = res.fgbox.copy()
        out[f'{name}_fgsum'] = np.array([int(res.foreground(k)[1].sum()) if res.status[k] < 2 else -1 for k in range(K)])
        out[f'{name}_ms'] = np.array(ts) * 1e3
        st = eng.solver_stats(); tot = max(st['cyc_total'], 1); wtot = max(st['cyc_wtotal'], 1)
        print(name, K, 'candidates', 'ms', [round(t * 1e3, 1) for t in ts], 'nf', int(res.info[:, 6].sum()), 'nh', int(res.info[:, 7].sum()),
              'ctrl busy %.3f eval %.3f hrows %.3f widle %.3f' % (st['cyc_ctrl'] / tot, st['cyc_eval'] / wtot, st['cyc_hrows'] / wtot, st['cyc_widle'] / wtot),
              {k: st[k] for k in ('n_eval', 'n_hrows', 'n_ctrl')},
              'skipped/unused/merged/hwasted', eng.eval_counts().sum(axis=0).tolist() if hasattr(eng.lib, 'sdsm_eval_counts') else None, 'it', res.info[:, 1:4].clip(0).sum(axis=0).tolist(), flush=True)
    np.savez_compressed(path, **out)


def compare(a, b):
    A, B = np.load(a), np.load(b)
    ok = True
    for key in A.files:
        if key.endswith('_ms'): print(key, A[key].round(1).tolist(), B[key].round(1).tolist()); continue
        x, y = A[key], B[key]
        same = x.shape == y.shape and np.array_equal(x.view(np.uint8), y.view(np.uint8))
        if not same:
            ok = False
            if x.shape == y.shape:
                d = np.nonzero((x != y).reshape(len(x), -1).any(axis=1))[0] if x.ndim > 1 else np.nonzero(x != y)[0]
                print(key, 'DIFFERS at', len(d), 'entries, first', d[:10].tolist())
            else: print(key, 'DIFFERS in shape', x.shape, y.shape)
        else: print(key, 'identical')
    print('BITWISE IDENTICAL' if ok else 'RESULTS DIFFER')
    return ok


if __name__ == '__main__':
    if sys.argv[1] == '--compare': sys.exit(0 if compare(sys.argv[2], sys.argv[3]) else 1)
    dump(sys.argv[1], sys.argv[2:] or ['gowt1'])
