"""Iterate traces of the pinned candidates of the headline workload (tests/test_gpu_full_parity.py::GOWT1_PINNED): device and
oracle interior-point iterates of the DSM solve side by side, the first iteration where they differ by more than 1e-6, and the
condition number of the KKT matrix z * scale / n * Hess psi there (TEST INFRASTRUCTURE: imports oracle/)."""
import sys
sys.path.insert(0, '.')
import numpy as np
import scipy.sparse
from oracle import dsm_oracle as orc
from oracle.make_full_golden import WORKLOAD_ARGS
from workloads.synth import make_workload
from superdsm_b200.objects import get_engine

ks = [int(v) for v in sys.argv[1:]] or [181, 202, 373]
w = make_workload(**WORKLOAD_ARGS['gowt1'])
cfg = dict(w['dsm_cfg'])
eng = get_engine(0)
eng.set_image(w['y'], w['y_mask'], w['atoms'], cfg['background_margin'])
eng.set_params(cfg)
res = eng.solve([w['footprints'][k] for k in ks])
np.set_printoptions(linewidth=200, precision=12)
for i, k in enumerate(ks):
    mask = orc.cvxprog_region_mask(w['y'], w['y_mask'], w['atoms'], w['footprints'][k], cfg['background_margin'])
    G = orc.smooth_matrix(mask, cfg['smooth_amount'], cfg['gaussian_shape_multiplier'], cfg['smooth_subsample'])
    n, m = int(mask.sum()), G.shape[1]
    scale_cp = cfg['scale'] / n
    J_ell = orc.Energy(w['y'], mask, cfg['epsilon'], cfg['alpha'], scipy.sparse.csr_matrix((n, 0)))
    J = orc.Energy(w['y'], mask, cfg['epsilon'], cfg['alpha'], G)
    theta = orc.elliptical_solution(J_ell, scale_cp)
    x0 = np.concatenate([theta, np.zeros(m)])
    tr, ev = [], []
    try: sol = orc._solve(J, x0, scale_cp, trace=tr, events=ev); o_status = sol['status']
    except Exception as e: o_status = f'raised {type(e).__name__}'
    st, nit, x, dtr = eng.trace_solve(i, x0, use_xi=True)
    ref = np.array([[t['f'], t['gap'], t['pres'], t['dres'], t['t'], t['s'], t['z'], t['relaxed']] for t in tr])
    print(f'== candidate {k}: n {n} m {m}; batch result: status {int(res.status[i])} energy {res.energy[i]:.10g} iterations {res.info[i, 1:4].tolist()} cp bits {int(res.info[i, 5])}')
    print(f'   oracle DSM solve: {o_status}, {len(tr) - 1} iterations, failed factorisations (iteration, relaxed counter): {[(e[1], e[2]) for e in ev]}; device trace: status {st}, {nit} iterations')
    first = None
    for it in range(min(len(ref), len(dtr))):
        rel = np.abs(dtr[it, :7] - ref[it, :7]) / np.maximum(np.abs(ref[it, :7]), 1e-12)
        if first is None and (rel.max() > 1e-6 or dtr[it, 7] != ref[it, 7]): first = it
    print(f'   iterates agree (1e-6 on f, gap, pres, dres, t, s, z; equal relaxed counters) up to iteration {first if first is not None else min(len(ref), len(dtr))}')
    lo = max(0, (first or 0) - 2)
    for it in range(lo, min(len(ref), len(dtr), (first or 0) + 3)):
        H = tr[it]['z'] * scale_cp * J.hessian(tr[it]['x'])
        try: cond = np.linalg.cond(H)
        except Exception: cond = np.nan
        print(f'   it {it:3d} relaxed dev/ora {int(dtr[it, 7]):2d}/{int(ref[it, 7]):2d}  f dev {dtr[it, 0]:.12g} ora {ref[it, 0]:.12g}  gap dev {dtr[it, 1]:.6g} ora {ref[it, 1]:.6g}  z dev {dtr[it, 6]:.6g} ora {ref[it, 6]:.6g}  cond(z scale/n Hess) {cond:.3g}')
    print(f'   final: J device {J(x):.10g}  J oracle {J(np.asarray(tr[-1]["x"])):.10g}  |grad| device {np.abs(J.grad(x)).max() * scale_cp:.3g} oracle {np.abs(J.grad(tr[-1]["x"])).max() * scale_cp:.3g} (scaled)')
