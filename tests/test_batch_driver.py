"""Batch mode: K image jobs in lock-step, their pending compute_objects calls merged into one launch."""
import os

import numpy as np
import pytest

from superdsm_b200.batch import LockstepBatch, object_rows, rasterize_labels, write_label_csv
from tests.compat.refshim import Image, Object

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')


def _stream_job(name, sink):
    """A job that replays the recorded GEM stream of a fixture: one compute_objects call per call of the reference's stage."""
    from oracle.make_gem_golden import GEM_ARGS
    from tests.test_gem_seam import _load
    from workloads.synth import make_workload
    z, fps, cover = _load(name)
    w = make_workload(**GEM_ARGS[name])
    y = Image(w['y'], w['y_mask'])

    def job(compute_objects):
        objs, pos = [], 0
        for n in z['call_len']:
            batch = [Object(fp) for fp in fps[pos:pos + n]]
            compute_objects(batch, y, w['atoms'], w['dsm_cfg'], None)
            objs += batch
            pos += n
        sink[name] = objs
        return len(objs)
    return job, z, fps, cover, w


def test_lockstep_rendezvous_with_a_fake_solver():
    """Host logic only: three jobs with 4, 2 and 7 calls; every launch merges the calls of all live jobs; results land in the right
    objects; an exception in the solver reaches the job that waits for it."""
    seen = []

    def solve(calls):
        seen.append(sorted(c.job for c in calls))
        for c in calls:
            for o in c.objects:
                st = Object(o.footprint)
                st.energy, st.on_boundary, st.is_optimal, st.processing_time = float(sum(o.footprint)) + 100 * c.job, False, True, 0.0
                st.fg_offset, st.fg_fragment = np.array([c.job, 0]), np.ones((1, len(o.footprint)), bool)
                o.set(st)

    def make(j, n_calls):
        def job(compute_objects):
            total = 0.0
            for k in range(n_calls):
                objs = [Object({k + 1, k + 2}), Object({k + 5})]
                compute_objects(objs, None, None, dict(scale=1.0, alpha=j % 2), None)        # jobs 0 and 2 share their hyper-parameters
                assert all(o.energy == sum(o.footprint) + 100 * j for o in objs)
                total += sum(o.energy for o in objs)
            return total
        return job
    lb = LockstepBatch(solve=solve)
    out = lb.run([make(0, 4), make(1, 2), make(2, 7)])
    assert out[0] == sum((k + 1) + (k + 2) + (k + 5) for k in range(4)) and len(out) == 3
    assert lb.calls == 13 and lb.launches == 7 + 2                      # jobs 0 and 2 merge (7 rounds), job 1 has its own hyper-parameters
    assert seen[0] in ([0, 2], [1]) and [2] in seen                      # the longest job ends alone

    def failing(calls): raise RuntimeError('device lost')
    with pytest.raises(RuntimeError, match='device lost'):
        LockstepBatch(solve=failing).run([make(0, 1)])


def test_result_formats(tmp_path):
    a, b = Object({1}), Object({2})
    a.fg_offset, a.fg_fragment = np.array([2, 3]), np.array([[1, 1], [0, 1]], bool)
    b.fg_offset, b.fg_fragment = np.array([0, 0]), np.array([[1]], bool)
    rows = object_rows([a, b])
    assert rows == [(1, 0.0, 0.0), (3, 3.7, 2.3)]
    write_label_csv(tmp_path / 'x.csv', rows)
    assert open(tmp_path / 'x.csv').read().splitlines() == ['"Object size","Center X","Center Y"', '"1","0.0","0.0"', '"3","3.7","2.3"']
    lab = rasterize_labels((5, 6), [a, b])
    assert lab.dtype == np.uint16 and lab[2, 3] == 1 and lab[3, 3] == 0 and lab[3, 4] == 1 and lab[0, 0] == 2 and (lab > 0).sum() == 4
    c = Object({3}); c.fg_offset, c.fg_fragment = np.array([2, 4]), np.array([[1, 1]], bool)          # overlaps a at (2, 4)
    lab = rasterize_labels((5, 6), [a, b, c])
    assert lab[2, 5] == 3 and lab[2, 4] in (1, 3) and (lab > 0).sum() == 5


@pytest.mark.skipif(not os.path.isdir('/root/reference/superdsm'), reason='the reference is only present in the build container')
def test_reference_gem_stage_as_a_lockstep_job():
    """The reference's UNMODIFIED GEM stage as one job and a recorded stream as another, solved by the oracle: the stage's cover
    costs equal the fixture's, and the two jobs' calls were merged."""
    from oracle import dsm_oracle as orc, ref_gem_harness
    from oracle.make_gem_golden import BETA, GEM_ARGS
    from workloads.synth import make_workload
    ref = ref_gem_harness.load()
    w = make_workload(**GEM_ARGS['toy'])

    def solve(calls):
        for c in calls:
            for o in c.objects:
                r = orc.compute_object(c.y.model, c.y.mask, c.atoms, o.footprint, c.cfg)
                o.energy, o.on_boundary, o.is_optimal, o.processing_time = r['energy'], r['on_boundary'], r['is_optimal'], 0.0
                o.fg_offset, o.fg_fragment = r['fg_offset'], r['fg_fragment']
    sink = {}
    replay, z, fps, cover, _ = _stream_job('toy', sink)

    def stage(compute_objects):
        return ref_gem_harness.run_gem(ref, w, lambda objs, y, atoms, cfg: compute_objects(objs, y, atoms, cfg, None), beta=BETA['toy'])
    lb = LockstepBatch(solve=solve)
    res, n = lb.run([stage, replay])
    assert res['cover'].costs == pytest.approx(float(z['cover_costs']), rel=1e-12)
    nonempty = int((z['call_len'] > 0).sum())                           # (the stage's last call has no candidates left)
    assert n == len(fps) and lb.launches == nonempty and lb.calls == 2 * nonempty
    assert np.allclose([o.energy for o in sink['toy']], z['energy'], rtol=1e-12, atol=1e-12)


def _mask_of(o, shape):
    m = np.zeros(shape, bool)
    r, c = (int(v) for v in o.fg_offset)
    h, w = o.fg_fragment.shape
    m[r:r + h, c:c + w] = o.fg_fragment
    return m


def _split_job(seed, parts, sink):
    """A job on another GOWT1-shaped image (same hyper-parameters): its candidate list in `parts` consecutive calls."""
    from workloads.synth import make_workload
    w = make_workload('gowt1', oversegment=2, max_set_size=2, seed=seed)
    y = Image(w['y'], w['y_mask'])
    chunks = np.array_split(np.arange(len(w['footprints'])), parts)

    def job(compute_objects):
        objs = []
        for idx in chunks:
            batch = [Object(w['footprints'][i]) for i in idx]
            compute_objects(batch, y, w['atoms'], w['dsm_cfg'], None)
            objs += batch
        sink[seed] = objs
        return len(objs)
    return job


@pytest.mark.gpu
def test_gem_stream_and_other_images_in_lockstep_on_the_device():
    """The recorded GEM stream of the GOWT1-shaped image (11 non-empty calls) and two other images of the batch (3 calls each) as
    lock-step jobs: 11 launches instead of 17, every result bit-identical to the job run alone, and the cover's (size, centroid)
    table equal to the fixture (the label-map format of tests/regression)."""
    from superdsm_b200.objects import compute_objects
    sink, alone = {}, {}
    job, z, fps, cover, w = _stream_job('gowt1', sink)
    lb = LockstepBatch()
    counts = lb.run([job, _split_job(2, 3, sink), _split_job(3, 3, sink)])
    nonempty = int((z['call_len'] > 0).sum())
    assert counts[0] == len(fps) and lb.launches == nonempty and lb.calls == nonempty + 6
    _stream_job('gowt1', alone)[0](compute_objects)                        # the same jobs through the single-image seam
    _split_job(2, 3, alone)(compute_objects)
    _split_job(3, 3, alone)(compute_objects)
    for key in ('gowt1', 2, 3):
        assert len(sink[key]) == len(alone[key])
        for a, b in zip(sink[key], alone[key]):
            assert a.energy == b.energy and a.is_optimal == b.is_optimal and a.on_boundary == b.on_boundary
            assert np.array_equal(a.fg_offset, b.fg_offset) and np.array_equal(a.fg_fragment, b.fg_fragment)
    by_fp = {tuple(sorted(o.footprint)): o for o in sink['gowt1']}
    rows = object_rows([by_fp[fp] for fp in cover])
    assert sorted(rows) == sorted(tuple(r) for r in z['stats'].tolist())
    lab = rasterize_labels(w['y'].shape, [by_fp[fp] for fp in cover])
    union = np.zeros(w['y'].shape, bool)
    for fp in cover: union |= _mask_of(by_fp[fp], union.shape)
    assert len(np.unique(lab)) - 1 == len(rows) and np.array_equal(lab > 0, union)       # every claimed pixel labelled, covers may overlap
