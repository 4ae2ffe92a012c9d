"""The seam under the reference's own callers (stage seam): ``tests/golden/gem_*.npz`` hold the candidate
stream, the energy thresholds and the final min-weight set cover of the UNMODIFIED ``GlobalEnergyMinimization`` code of the
reference driven through the oracle seam (``oracle/make_gem_golden.py``).  On the device the same stream goes through
``superdsm_b200.objects.compute_objects`` call by call: energies, pruning decisions, the cover and the (size, centroid)
multiset of the cover's objects (the label-map comparison of ``tests/regression/validate.py``) must come out the same."""
import os

import numpy as np
import pytest

from tests.compat import setcover

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')


def _load(name):
    z = np.load(os.path.join(GOLDEN, f'gem_{name}.npz'))
    offs = np.concatenate([[0], np.cumsum(z['fp_len'])])
    fps = [tuple(int(a) for a in z['fp_flat'][offs[k]:offs[k + 1]]) for k in range(len(z['fp_len']))]
    coffs = np.concatenate([[0], np.cumsum(z['cover_len'])])
    cover = sorted(tuple(int(a) for a in z['cover_flat'][coffs[k]:coffs[k + 1]]) for k in range(len(z['cover_len'])))
    return z, fps, cover


def _family(z, fps, energies):
    """The family the cover sees: atoms, universes, then per generation only the sets that beat their threshold
    (globalenergymin.py:357-364), in stream order; `stages` for setcover.cover_by_cluster."""
    keep_fps, keep_e, stages = [], [], []
    pos = 0
    for n in z['call_len']:
        for k in range(pos, pos + n):
            t = z['threshold'][k]
            if np.isnan(t) or energies[k] < t:
                keep_fps.append(fps[k]); keep_e.append(float(energies[k]))
        stages.append(len(keep_fps))
        pos += n
    return keep_fps, keep_e, stages


@pytest.mark.parametrize('name', ['toy', 'bbbc039', 'gowt1'])
def test_restated_set_cover_reproduces_the_reference_cover(name):
    """CPU: tests/compat/setcover.py on the recorded oracle energies gives the cover the reference's MinSetCover gave."""
    if not os.path.isfile(os.path.join(GOLDEN, f'gem_{name}.npz')): pytest.skip('fixture not generated')
    z, fps, cover = _load(name)
    keep_fps, keep_e, stages = _family(z, fps, z['energy'])
    got = setcover.cover_by_cluster(keep_fps, keep_e, z['atom_cluster'], float(z['beta']), stages)
    assert got == cover
    assert sum(z['generation_sizes'][1:]) == len(keep_fps) - stages[1]     # the accepted sets are the stage's generations


@pytest.mark.gpu
@pytest.mark.parametrize('name', ['toy', 'bbbc039', 'gowt1'])
def test_gem_stream_through_the_gpu_seam(name):
    if not os.path.isfile(os.path.join(GOLDEN, f'gem_{name}.npz')): pytest.skip('fixture not generated')
    from oracle.make_full_golden import mask_hash
    from oracle.make_gem_golden import GEM_ARGS, object_stats
    from superdsm_b200.objects import compute_objects
    from tests.compat.refshim import Image, Object
    from workloads.synth import make_workload
    z, fps, cover = _load(name)
    w = make_workload(**GEM_ARGS[name])
    assert float(w['y'].sum()) == float(z['image_checksum'])
    y = Image(w['y'], w['y_mask'])
    scale = float(w['dsm_cfg']['scale'])
    count = np.bincount(w['atoms'].reshape(-1))
    objs, pos = [], 0
    for n in z['call_len']:                                   # one compute_objects call per call of the stage
        batch = [Object(fp) for fp in fps[pos:pos + n]]
        compute_objects(batch, y, w['atoms'], w['dsm_cfg'], None)
        objs += batch
        pos += n
    energy = np.array([o.energy for o in objs])
    npix = np.array([sum(count[a] for a in fp) for fp in fps], float)
    tol = 1e-6 * np.abs(z['energy']) + 2e-7 * npix / scale
    bad = np.nonzero(np.abs(energy - z['energy']) > tol)[0]
    # (irregular, nearly separable fits may stop elsewhere: tests/test_gpu_full_parity.py; what matters here are the decisions below)
    assert len(bad) <= 0.03 * len(fps), (len(bad), bad[:20], energy[bad][:5], z["energy"][bad][:5])
    # pruning decisions of the stage (energy < threshold), candidate by candidate
    has_t = ~np.isnan(z['threshold'])
    assert np.array_equal(energy[has_t] < z['threshold'][has_t], z['energy'][has_t] < z['threshold'][has_t])
    # the cover from the device's energies is the reference's cover, and its objects have the same foregrounds
    keep_fps, keep_e, stages = _family(z, fps, energy)
    got = setcover.cover_by_cluster(keep_fps, keep_e, z['atom_cluster'], float(z['beta']), stages)
    assert got == cover
    by_fp = {tuple(sorted(o.footprint)): o for o in objs}
    sol = [by_fp[fp] for fp in cover]
    stats = np.array(object_stats(sol), float)
    assert stats.shape == z['stats'].shape and np.array_equal(stats, z['stats'])       # identical label map (size / centroid multiset)
    hashes = {fp: h for fp, h in zip(fps, z['mask_hash'])}
    assert all(mask_hash(o.fg_offset, o.fg_fragment) == hashes[fp] for fp, o in zip(cover, sol))
