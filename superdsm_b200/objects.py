"""Drop-in for the hot path of the reference's ``superdsm/objects.py``.

Same names, arguments and result fields as the reference; the per-candidate Python/Ray/cvxopt/MKL work of
``objects.py:177-284,361-412`` is replaced by one batched call into the CUDA library.

    compute_objects(objects, y, atoms, dsm_cfg, log_root_dir, status_line, out)      objects.py:243-267
    _compute_objects(objects, y, atoms, x_map, lock, dsm_cfg, log_root_dir)          objects.py:270-284  (the seam)
    cvxprog(region, scale, epsilon, alpha, smooth_amount, ...)                       objects.py:361-412

``INTEGRATION.md`` shows how a maintainer binds these into the reference.
"""
import time

import numpy as np

from . import _lib
from .dsm import DeformableShapeModel, Energy
from .engine import Engine, SolverError

_engines = {}


def get_engine(device=None):
    """The process-wide engine of a device (one process per GPU: LOCAL_RANK picks the default device)."""
    import os
    if device is None: device = int(os.environ.get('LOCAL_RANK', 0))
    if device not in _engines: _engines[device] = Engine(device)
    return _engines[device]


class BaseObject:
    """Segmentation mask = foreground fragment + offset (reference ``objects.py:12-50``)."""

    def __init__(self):
        self.fg_offset = None
        self.fg_fragment = None

    def fill_foreground(self, out, value=True):
        assert self.fg_offset is not None
        assert self.fg_fragment is not None
        sel = np.s_[self.fg_offset[0]:self.fg_offset[0] + self.fg_fragment.shape[0],
                    self.fg_offset[1]:self.fg_offset[1] + self.fg_fragment.shape[1]]
        out[sel] = value * self.fg_fragment
        return sel


class Object(BaseObject):
    """A set of atomic image regions and, once computed, its fitted segmentation (reference ``objects.py:53-145``)."""

    def __init__(self):
        self.footprint = set()
        self.energy = np.nan
        self.on_boundary = np.nan
        self.is_optimal = np.nan
        self.processing_time = np.nan

    def get_mask(self, atoms):
        return np.isin(atoms, list(self.footprint)).reshape(atoms.shape)

    def get_cvxprog_region(self, y, atoms, background_margin):
        import scipy.ndimage as ndi
        region = y.get_region(self.get_mask(atoms))
        region.mask = np.logical_and(region.mask, ndi.distance_transform_edt(y.model <= 0) <= background_margin)
        return region

    def set(self, state):
        self.fg_fragment = state.fg_fragment.copy() if state.fg_fragment is not None else None
        self.fg_offset = state.fg_offset.copy() if state.fg_offset is not None else None
        self.footprint = set(state.footprint)
        self.energy = state.energy
        self.on_boundary = state.on_boundary
        self.is_optimal = state.is_optimal
        self.processing_time = state.processing_time
        return self

    def copy(self):
        return Object().set(self)


def extract_foreground_fragment(fg_mask):
    """Minimal rectangle of the foreground and its offset (reference ``objects.py:148-174``)."""
    if fg_mask.any():
        rows, cols = fg_mask.any(axis=1), fg_mask.any(axis=0)
        rmin, rmax = np.where(rows)[0][[0, -1]]
        cmin, cmax = np.where(cols)[0][[0, -1]]
        return np.array([rmin, cmin]), fg_mask[rmin:rmax + 1, cmin:cmax + 1]
    return np.zeros(2, int), np.zeros((1, 1), bool)


class CvxprogError(Exception):
    def __init__(self, *args, cidx=None, cause=None):
        super().__init__(*args)
        self.cidx = cidx

    def __str__(self):
        messages = [str(arg) for arg in self.args]
        if self.cidx is not None: messages.append(f'cidx: {self.cidx}')
        return ', '.join(messages)


DEFAULT_COMPUTING_STATUS_LINE = ('Computing objects', 'Computed objects')

_image_cache = {}


def _bind_image(engine, y, atoms, background_margin):
    """Uploads (y, atoms) unless the engine already holds exactly these arrays for this margin."""
    key = (id(y.model), id(y.mask), id(atoms), y.model.shape, float(background_margin),
           float(y.model.reshape(-1)[::997].sum()), int(np.asarray(atoms).reshape(-1)[::997].sum()))
    if _image_cache.get(id(engine)) != key:
        engine.set_image(y.model, y.mask, atoms, background_margin)
        _image_cache[id(engine)] = key


def _result_to_object(res, k, obj, dt):
    st = int(res.status[k])
    if st == _lib.ST_ERROR: raise CvxprogError('elliptical initialisation failed', cidx=k)
    if st == _lib.ST_BADREGION: raise SolverError(f'candidate {k}: region mismatch between host and device (flags in info)')
    state = Object()
    state.footprint = set(obj.footprint)
    if st == _lib.ST_SKIPPED:                                   # objects.py:184-191
        state.fg_offset, state.fg_fragment = np.zeros(2, int), np.zeros((1, 1), bool)
        state.energy, state.on_boundary, state.is_optimal, state.processing_time = 0., False, False, 0
        return state, False
    state.fg_offset, state.fg_fragment = res.foreground(k)
    state.energy = float(res.energy[k])
    state.on_boundary = bool(res.on_boundary[k])
    state.is_optimal = (st == _lib.ST_OPTIMAL)
    state.processing_time = dt
    return state, st == _lib.ST_FALLBACK


def _compute_objects(objects, y, atoms, x_map, lock, dsm_cfg, log_root_dir):
    """Yields ``(cidx, Object, fallback)`` for every object, like the reference's generator (``objects.py:270-284``).

    ``x_map``, ``lock`` and ``log_root_dir`` are accepted for signature compatibility; the padded coordinate map is not
    needed (the result is rasterised on the device), there is no host-side matrix allocation to lock, and there are no
    per-candidate solver logs.
    """
    objects = list(objects)
    if len(objects) == 0: return
    cfg = dict(dsm_cfg)
    cfg.pop('smooth_mat_max_allocations', None)
    engine = get_engine()
    _bind_image(engine, y, atoms, cfg['background_margin'])
    engine.set_params(cfg)
    t0 = time.time()
    res = engine.solve([obj.footprint for obj in objects], borrow=True)
    dt = (time.time() - t0) / len(objects)
    # all states are built before the first one is handed out: the masks are views of the engine's staging buffer
    states = [_result_to_object(res, cidx, obj, dt) for cidx, obj in enumerate(objects)]
    for cidx, (state, fallback) in enumerate(states):
        yield cidx, state, fallback


_compute_objects._DEBUG = False


def compute_objects(objects, y, atoms, dsm_cfg, log_root_dir, status_line=DEFAULT_COMPUTING_STATUS_LINE, out=None):
    """Computes ``energy``, ``on_boundary``, ``is_optimal``, ``processing_time``, ``fg_fragment``, ``fg_offset`` of all
    objects in place (reference ``objects.py:243-267``)."""
    dsm_cfg = dict(dsm_cfg)
    dsm_cfg.pop('smooth_mat_max_allocations', None)
    objects = list(objects)
    fallbacks = 0
    for ret in _compute_objects(objects, y, atoms, None, None, dsm_cfg, log_root_dir):
        objects[ret[0]].set(ret[1])
        if ret[2]: fallbacks += 1
    if out is not None and hasattr(out, 'write'):
        out.write(f'{status_line[1]}: {len(objects)} ({fallbacks}x fallback)')


def cvxprog(region, scale, epsilon, alpha, smooth_amount, smooth_subsample, gaussian_shape_multiplier, smooth_mat_allocation_lock=None,
            smooth_mat_dtype='float32', sparsity_tol=0, hessian_sparsity_tol=0, init=None, cachesize=0, cachetest=None, cp_timeout=None):
    """Fits a deformable shape model to one image region (reference ``objects.py:361-412``).

    Returns ``(J, DeformableShapeModel, status)`` with ``status`` in ``('optimal', 'fallback')``.
    """
    if hessian_sparsity_tol != 0: raise SolverError('hessian_sparsity_tol != 0 is not implemented')
    engine = get_engine()
    mask = np.asarray(region.mask, dtype=bool)
    # a one-atom image whose usable pixels are exactly the region (margin = inf keeps every masked pixel)
    engine.set_image(region.model, mask, mask.astype(np.int32), np.inf)
    _image_cache.pop(id(engine), None)
    engine.set_params(dict(scale=scale, epsilon=epsilon, alpha=alpha, smooth_amount=smooth_amount, smooth_subsample=smooth_subsample,
                           gaussian_shape_multiplier=gaussian_shape_multiplier, smooth_mat_dtype=smooth_mat_dtype,
                           sparsity_tol=sparsity_tol, init=init))
    res = engine.solve([{1}])
    st = int(res.status[0])
    if st == _lib.ST_ERROR: raise CvxprogError('elliptical initialisation failed')
    if st == _lib.ST_BADREGION: raise SolverError('region mismatch between host and device')
    if st == _lib.ST_SKIPPED: raise SolverError('cvxprog on a region with a single foreground pixel (callers skip these, objects.py:184)')
    params = res.params(0)
    J = Energy(engine, 0, int(mask.sum()), res.m(0), params, res.energy[0])
    return J, DeformableShapeModel(params), 'optimal' if st == _lib.ST_OPTIMAL else 'fallback'
