"""Loads the reference's GEM stage UNCHANGED (``superdsm/globalenergymin.py``, ``minsetcover.py``, ``maxsetpack.py``, ``atoms.py``,
``pipeline.py``) on top of ``oracle/ref_harness.py``, so that its candidate enumeration, pruning and set cover can be driven
with either seam: the oracle's ``compute_objects`` or ``superdsm_b200.objects.compute_objects``.

TEST INFRASTRUCTURE, build container only (``/root/reference`` must exist).  Additional stubs: ``cvxpy`` (imported by
``maxsetpack.py:4``, unused), ``skimage.morphology.disk`` / ``binary_dilation`` and ``skimage.segmentation`` (``atoms.py:4-5``;
scikit-image is not installed: scipy.ndimage equivalents).
"""
import importlib.util
import os
import sys
import types
import warnings

import numpy as np
import scipy.ndimage as ndi

from . import ref_harness


def load():
    ref = ref_harness.load()
    pkg = sys.modules['superdsm']
    if getattr(pkg, '_gem_loaded', False):
        return types.SimpleNamespace(gem=pkg.globalenergymin, atoms=pkg.atoms, objects=pkg.objects, image=pkg.image, minsetcover=pkg.minsetcover)

    def disk(radius):
        r = np.arange(-radius, radius + 1)
        return (r[:, None] ** 2 + r[None, :] ** 2 <= radius ** 2).astype(np.uint8)
    morph = sys.modules['skimage.morphology']
    morph.disk = disk
    morph.binary_dilation = lambda img, footprint=None: ndi.binary_dilation(img, structure=footprint)
    seg = types.ModuleType('skimage.segmentation')
    sys.modules['skimage.segmentation'] = seg
    sys.modules['skimage'].segmentation = seg
    sys.modules['cvxpy'] = types.ModuleType('cvxpy')

    def load_file(name):
        path = os.path.join(ref_harness.REFERENCE_ROOT, 'superdsm', name + '.py')
        spec = importlib.util.spec_from_file_location('superdsm.' + name, path)
        mod = importlib.util.module_from_spec(spec)
        sys.modules['superdsm.' + name] = mod
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            spec.loader.exec_module(mod)
        setattr(pkg, name, mod)
        return mod

    for name in ('pipeline', 'atoms', 'minsetcover', 'maxsetpack', 'globalenergymin'): load_file(name)
    pkg._gem_loaded = True
    return types.SimpleNamespace(gem=pkg.globalenergymin, atoms=pkg.atoms, objects=pkg.objects, image=pkg.image, minsetcover=pkg.minsetcover)


def run_gem(ref, w, compute_objects, beta=0.0, pruning='exact', max_seed_distance=np.inf, record=None):
    """The reference's ``_compute_generations`` (globalenergymin.py:183-271) on workload `w` (workloads.synth) with the given
    seam.  ``record``: list that receives one dict per ``compute_objects`` call (footprints, and after the call the results)."""
    y_img = ref.image.Image.create_from_array(w['y'], normalize=False, mask=w['y_mask'])
    adj = ref.atoms.AtomAdjacencyGraph(w['atoms'], w['clusters'], w['y'] > 0, w['atom_seeds'], 'muted')
    calls = record if record is not None else []

    def seam(objects, y, atoms, dsm_cfg, log_root_dir, status_line=None, out=None):
        objects = list(objects)
        compute_objects(objects, y, atoms, dsm_cfg)
        calls.append(dict(footprints=[sorted(o.footprint) for o in objects], energy=[o.energy for o in objects],
                          is_optimal=[bool(o.is_optimal) for o in objects], objects=objects))
    saved = ref.gem.compute_objects
    ref.gem.compute_objects = seam
    try:
        generations, costs, cover, objects, performance = ref.gem._compute_generations(
            adj, y_img, w['atoms'], None, pruning, dict(w['dsm_cfg']), beta, ref.gem.DEFAULT_MAX_ITER, ref.gem.DEFAULT_GAMMA,
            max_seed_distance, ref.gem.DEFAULT_MAX_WORK_AMOUNT, 'muted')
    finally:
        ref.gem.compute_objects = saved
    return dict(generations=generations, costs=costs, cover=cover, objects=objects, performance=performance, calls=calls, adjacencies=adj)
