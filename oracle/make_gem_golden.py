"""Generates ``tests/golden/gem_<workload>.npz``: the REFERENCE's own GEM stage (``superdsm/globalenergymin.py`` unchanged, via
``oracle/ref_gem_harness.py``) driven through the oracle seam on a BASELINE-shaped synthetic image.

TEST INFRASTRUCTURE, build container only:  ``python -m oracle.make_gem_golden bbbc039 [processes]``.  The fixture records
* the candidate stream: the footprints of every ``compute_objects`` call the stage made (atoms, universes, generations), with
  the oracle's energy / is_optimal / foreground hash of every candidate and the energy threshold the stage compared it with
  (``globalenergymin.py:360-364``; nan for atoms and universes),
* the min-weight set cover the stage ended with (footprints), its costs, and the (size, centroid) multiset of the cover's
  objects -- the format ``tests/regression/validate.py:26-100`` compares label maps in.
"""
import os
import sys

for _v in ('OMP_NUM_THREADS', 'OPENBLAS_NUM_THREADS', 'MKL_NUM_THREADS'): os.environ[_v] = '1'

import numpy as np                                   # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path: sys.path.insert(0, ROOT)

from oracle.make_full_golden import WORKLOAD_ARGS, mask_hash     # noqa: E402

OUT = os.path.join(ROOT, 'tests', 'golden')
#: workload -> arguments of workloads.synth.make_workload (only image and atoms matter: the stage enumerates the candidates itself)
GEM_ARGS = {'toy': dict(name='toy', oversegment=2, max_set_size=1), 'gowt1': dict(WORKLOAD_ARGS['gowt1']),
            'bbbc039': dict(name='bbbc039', oversegment=2, max_set_size=1, seed=39)}
#: workload -> beta of the GEM stage (globalenergymin.py:144; AF_beta x scale^2 under automation)
BETA = {'bbbc039': 0.0, 'gowt1': 1188.0, 'toy': 150.0}
_job = None


def _fit(fp):
    from oracle import dsm_oracle as orc
    y, y_mask, atoms, cfg = _job
    try:
        r = orc.compute_object(y, y_mask, atoms, fp, cfg)
    except orc.CvxprogError:
        return None
    return r['energy'], r['on_boundary'], r['is_optimal'], r['fg_offset'], r['fg_fragment']


def object_stats(objects):
    """(size, centroid x, centroid y) of every object's foreground, as tests/regression/validate.py:33-38 lists them."""
    rows = []
    for o in objects:
        frag = np.asarray(o.fg_fragment, bool)
        if not frag.any(): continue
        rr, cc = np.nonzero(frag)
        rows.append((int(frag.sum()), round(float(cc.mean() + o.fg_offset[1]), 1), round(float(rr.mean() + o.fg_offset[0]), 1)))
    return sorted(rows)


def main():
    import multiprocessing as mp
    global _job
    from oracle import ref_gem_harness
    from workloads.synth import make_workload
    wname = sys.argv[1]
    procs = int(sys.argv[2]) if len(sys.argv) > 2 else (os.cpu_count() or 1)
    w = make_workload(**GEM_ARGS[wname])
    ref = ref_gem_harness.load()
    _job = (w['y'], w['y_mask'], w['atoms'], w['dsm_cfg'])
    pool = mp.get_context('fork').Pool(procs)
    thresholds = []

    def compute_objects(objects, y, atoms, dsm_cfg):
        caller = sys._getframe(2).f_locals                         # _process_generation's frame (None for atoms / universes)
        thr = caller.get('new_objects_energy_thresholds')
        thresholds.append(list(thr) if thr is not None else [np.nan] * len(objects))
        for o, r in zip(objects, pool.map(_fit, [sorted(o.footprint) for o in objects], chunksize=2)):
            if r is None: raise RuntimeError('CvxprogError in the GEM stream')
            o.energy, o.on_boundary, o.is_optimal, o.fg_offset, o.fg_fragment = r
            o.processing_time = 0.0
    res = ref_gem_harness.run_gem(ref, w, compute_objects, beta=BETA.get(wname, 0.0))
    pool.close()
    calls = res['calls']
    fp_flat, fp_len, call_len, energy, optimal, hashes, thr = [], [], [], [], [], [], []
    for call, t in zip(calls, thresholds):
        call_len.append(len(call['footprints']))
        for fp, o, tv in zip(call['footprints'], call['objects'], t):
            fp_flat += fp; fp_len.append(len(fp)); energy.append(o.energy); optimal.append(bool(o.is_optimal)); thr.append(tv)
            hashes.append(mask_hash(o.fg_offset, o.fg_fragment) if o.fg_fragment is not None else 0)
    cover = res['cover']
    sol = cover.solution
    adj = res['adjacencies']
    atom_cluster = np.array([adj.get_cluster_label(a) for a in sorted(adj.atom_labels)])
    stats = object_stats(sol)
    path = os.path.join(OUT, f'gem_{wname}.npz')
    np.savez_compressed(path, fp_flat=np.array(fp_flat, np.int32), fp_len=np.array(fp_len, np.int32), call_len=np.array(call_len, np.int32),
                        energy=np.array(energy), is_optimal=np.array(optimal), mask_hash=np.array(hashes, np.uint64), threshold=np.array(thr, float),
                        cover_flat=np.array([a for o in sol for a in sorted(o.footprint)], np.int32), cover_len=np.array([len(o.footprint) for o in sol], np.int32),
                        cover_costs=np.float64(cover.costs), beta=np.float64(cover.beta), atom_cluster=atom_cluster, stats=np.array(stats, float),
                        image_checksum=np.float64(w['y'].sum()), atoms_checksum=np.int64(w['atoms'].astype(np.int64).sum()),
                        generation_sizes=np.array([len(g) for g in res['generations']], np.int32))
    perf = res['performance']
    print(f'wrote {path}: {len(calls)} compute_objects calls, {len(energy)} candidates (bruteforce {perf.overall_object_count}), '
          f'cover {len(sol)} objects, costs {cover.costs:.6f}, generations {[len(g) for g in res["generations"]]}, {os.path.getsize(path) // 1024} KiB')


if __name__ == '__main__':
    main()
