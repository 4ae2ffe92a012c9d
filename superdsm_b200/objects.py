"""Drop-in for the hot path of the reference's ``superdsm/objects.py``.

Same names, arguments and result fields as the reference; the per-candidate Python/Ray/cvxopt/MKL work of
``objects.py:177-284,361-412`` is replaced by batched calls into the CUDA library.

    compute_objects(objects, y, atoms, dsm_cfg, log_root_dir, status_line, out)      objects.py:243-267
    _compute_objects(objects, y, atoms, x_map, lock, dsm_cfg, log_root_dir)          objects.py:270-284  (the seam)
    cvxprog(region, scale, epsilon, alpha, smooth_amount, ...)                       objects.py:361-412
    compute_objects_multi(jobs, dsm_cfg)                                             several images in one launch (batch mode)

The functions work on the caller's own classes: an image is anything with ``.model`` (float64 H x W) and ``.mask``, an
object anything with ``.footprint`` and ``.set(state)``; the states handed to ``set`` carry the fields the reference's
``Object.set`` copies (``objects.py:130-140``).  ``INTEGRATION.md`` shows how a maintainer binds these into the reference.
"""
import hashlib
import time

import numpy as np

from . import _lib
from .dsm import Energy, ModelParams
from .engine import Engine, SolverError

_engines = {}


def get_engine(device=None):
    """The process-wide engine of a device (one process per GPU: LOCAL_RANK picks the default device)."""
    import os
    if device is None: device = int(os.environ.get('LOCAL_RANK', 0))
    if device not in _engines: _engines[device] = Engine(device)
    return _engines[device]


class CvxprogError(Exception):
    """Both elliptical passes failed (the reference raises its own ``CvxprogError``, ``objects.py:351-353``)."""

    def __init__(self, *args, cidx=None):
        super().__init__(*args)
        self.cidx = cidx

    def __str__(self):
        return ', '.join([str(a) for a in self.args] + ([f'cidx: {self.cidx}'] if self.cidx is not None else []))


class ObjectState(object):
    """Result of one candidate: the fields ``Object.set`` reads (``objects.py:130-140``)."""
    __slots__ = ('footprint', 'energy', 'on_boundary', 'is_optimal', 'processing_time', 'fg_offset', 'fg_fragment')


DEFAULT_COMPUTING_STATUS_LINE = ('Computing objects', 'Computed objects')


def _digest(*arrays):
    h = hashlib.blake2b(digest_size=16)
    for a in arrays:
        a = np.ascontiguousarray(a)
        h.update(str((a.shape, a.dtype.str)).encode())
        h.update(a.view(np.uint8).reshape(-1))
    return h.digest()


_bound = {}


def _bind_image(engine, y, atoms, background_margin):
    """Uploads (y, atoms) unless the engine already holds exactly these contents for this margin (the stages call
    ``compute_objects`` once per generation on the same image): the key is a hash of the full contents, ~2 ms per megapixel."""
    key = (_digest(y.model, y.mask, atoms), float(background_margin))
    if _bound.get(id(engine)) != key or engine._bound != [engine._image]:
        engine.set_image(y.model, y.mask, atoms, background_margin)
        _bound[id(engine)] = key


def forget_images():
    """Drops the record of what is resident (the next call uploads again)."""
    _bound.clear()


def _state(res, k, footprint, dt):
    st = int(res.status[k])
    if st == _lib.ST_ERROR: raise CvxprogError('elliptical initialisation failed', cidx=k)
    if st == _lib.ST_BADREGION: raise SolverError(f'candidate {k}: region mismatch between host and device (flags in info)')
    state = ObjectState()
    state.footprint = set(footprint)
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


def _check_cfg(cfg):
    if not (cfg.get('hessian_sparsity_tol', 0) >= 0): raise SolverError('hessian_sparsity_tol must be >= 0')


def _compute_objects(objects, y, atoms, x_map, lock, dsm_cfg, log_root_dir):
    """Yields ``(cidx, state, fallback)`` for every object, like the reference's generator (``objects.py:270-284``).

    ``x_map``, ``lock`` and ``log_root_dir`` are accepted for signature compatibility; the padded coordinate map is not
    needed (the result is rasterised on the device), there is no host-side matrix allocation to lock, and there are no
    per-candidate solver logs.
    """
    objects = list(objects)
    if len(objects) == 0: return
    cfg = dict(dsm_cfg)
    cfg.pop('smooth_mat_max_allocations', None)
    _check_cfg(cfg)
    engine = get_engine()
    _bind_image(engine, y, atoms, cfg['background_margin'])
    engine.set_params(cfg, skip_single_fg=True)
    t0 = time.time()
    res = engine.solve([obj.footprint for obj in objects], borrow=True)
    dt = (time.time() - t0) / len(objects)
    # all states are built before the first one is handed out: the masks are views of the engine's staging buffer
    states = [_state(res, cidx, obj.footprint, dt) for cidx, obj in enumerate(objects)]
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
    return fallbacks


class ImageBatch(object):
    """Several images resident on one device at once (batch mode, ``batch.py:288-361``): the candidates of all of them go
    through ONE launch of the persistent solver, so that the long dependency chains of one image overlap the bulk of the
    others.  ``images``: a list of ``(y, atoms)`` as ``compute_objects`` takes them."""

    def __init__(self, images, background_margin, engine=None):
        self.engine = engine or get_engine()
        self.handles = [self.engine.upload_image(y.model, y.mask, atoms, background_margin) for y, atoms in images]

    def compute_objects(self, objects_per_image, dsm_cfg):
        """``objects_per_image[i]``: the objects of image i (may be empty).  Sets their fields in place; returns the number
        of fallbacks per image."""
        cfg = dict(dsm_cfg)
        cfg.pop('smooth_mat_max_allocations', None)
        _check_cfg(cfg)
        eng = self.engine
        eng.bind_images(self.handles)
        _bound.pop(id(eng), None)
        eng.set_params(cfg, skip_single_fg=True)
        lists = [list(objs) for objs in objects_per_image]
        t0 = time.time()
        res = eng.solve_multi([[o.footprint for o in objs] for objs in lists], borrow=True)
        dt = (time.time() - t0) / max(1, len(res))
        fallbacks, k = [0] * len(lists), 0
        states = []
        for i, objs in enumerate(lists):
            for o in objs:
                states.append((i, o, _state(res, k, o.footprint, dt)))
                k += 1
        for i, o, (state, fb) in states:
            o.set(state)
            fallbacks[i] += int(fb)
        return fallbacks


def compute_objects_multi(jobs, dsm_cfg):
    """``jobs``: list of ``(objects, y, atoms)``, one per image (at most 32).  All candidates of all images in one launch."""
    batch = ImageBatch([(y, atoms) for _, y, atoms in jobs], dict(dsm_cfg)['background_margin'])
    return batch.compute_objects([objs for objs, _, _ in jobs], dsm_cfg)


def cvxprog(region, scale, epsilon, alpha, smooth_amount, smooth_subsample, gaussian_shape_multiplier, smooth_mat_allocation_lock=None,
            smooth_mat_dtype='float32', sparsity_tol=0, hessian_sparsity_tol=0, init=None, cachesize=0, cachetest=None, cp_timeout=None):
    """Fits a deformable shape model to one image region (reference ``objects.py:361-412``).

    Returns ``(J, model, status)`` with ``status`` in ``('optimal', 'fallback')``; ``model.array`` holds (theta, xi).
    A region with a single foreground pixel is fitted like any other: the skip of such regions is a rule of
    ``_compute_object`` (``objects.py:184-191``), not of ``cvxprog``.
    """
    engine = get_engine()
    mask = np.asarray(region.mask, dtype=bool)
    # a one-atom image whose usable pixels are exactly the region (margin = inf keeps every masked pixel)
    engine.set_image(region.model, mask, mask.astype(np.int32), np.inf)
    _bound.pop(id(engine), None)
    engine.set_params(dict(scale=scale, epsilon=epsilon, alpha=alpha, smooth_amount=smooth_amount, smooth_subsample=smooth_subsample,
                           gaussian_shape_multiplier=gaussian_shape_multiplier, smooth_mat_dtype=smooth_mat_dtype,
                           sparsity_tol=sparsity_tol, hessian_sparsity_tol=hessian_sparsity_tol, init=init), skip_single_fg=False)
    res = engine.solve([{1}])
    st = int(res.status[0])
    if st == _lib.ST_ERROR: raise CvxprogError('elliptical initialisation failed')
    if st == _lib.ST_BADREGION: raise SolverError('region mismatch between host and device')
    params = res.params(0)
    J = Energy(engine, 0, int(mask.sum()), res.m(0), params, res.energy[0])
    return J, ModelParams(params), 'optimal' if st == _lib.ST_OPTIMAL else 'fallback'
