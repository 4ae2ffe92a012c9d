"""Replays the recorded GEM candidate streams (tests/golden/gem_*.npz) through the GPU seam and saves the energies for analysis."""
import sys
sys.path.insert(0, '.')
import numpy as np
from tests.test_gem_seam import _load
from oracle.make_gem_golden import GEM_ARGS
from superdsm_b200.objects import compute_objects, get_engine
from tests.compat.refshim import Image, Object
from workloads.synth import make_workload
for name in sys.argv[1:] or ['bbbc039', 'gowt1']:
    z, fps, cover = _load(name)
    w = make_workload(**GEM_ARGS[name])
    y = Image(w['y'], w['y_mask'])
    objs, pos = [], 0
    infos = []
    for n in z['call_len']:
        batch = [Object(fp) for fp in fps[pos:pos + n]]
        compute_objects(batch, y, w['atoms'], w['dsm_cfg'], None)
        res_info = get_engine()._last
        objs += batch; pos += n
    energy = np.array([o.energy for o in objs])
    opt = np.array([bool(o.is_optimal) for o in objs])
    np.savez_compressed(f'gpurun_out/gemreplay_{name}.npz', energy=energy, is_optimal=opt)
    has_t = ~np.isnan(z['threshold'])
    dd = (energy < z['threshold']) != (z['energy'] < z['threshold'])
    print(name, 'candidates', len(fps), 'decision flips', int((dd & has_t).sum()), np.nonzero(dd & has_t)[0][:20])
