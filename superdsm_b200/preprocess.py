"""The reference's ``Preprocessing`` stage on the device (``superdsm/preprocess.py:39-68``): the offset image
``y = G_sigma1 * g - offset``, where the offset blends two wide Gaussian filters of the raw image (one of them of the clipped
image) by the distance to the bright (clipped) areas.  Three separable Gaussian filters (sigma up to 40 and more: 321 taps)
and an exact Euclidean distance transform per image -- per-image work in front of the hot path.

The result is bit-identical to the scipy statement (``tests/test_gpu_round2.py``): the kernels sum in the order of scipy's
``correlate1d`` for symmetric kernels, with its 'reflect' boundary handling.  The two scalar reductions of the stage
(``g_raw.std()``, ``g_raw.mean()``) are numpy's (their pairwise summation order is part of the result); everything per pixel
runs on the device.
"""
import ctypes
import math

import numpy as np

from .engine import SolverError
from .objects import get_engine


def gaussian_weights(sigma, truncate=4.0):
    """The weights ``scipy.ndimage.gaussian_filter1d`` correlates with (``_gaussian_kernel1d`` reversed) and their radius."""
    sd = float(sigma)
    radius = int(truncate * sd + 0.5)
    x = np.arange(-radius, radius + 1)
    phi = np.exp(-0.5 / (sd * sd) * x ** 2)
    phi = phi / phi.sum()
    return np.ascontiguousarray(phi[::-1]), radius


def offset_image(g_raw, sigma1=math.sqrt(2), sigma2=40, offset_clip=3, lower_clip_mean=False, engine=None):
    """``Preprocessing.process`` (``preprocess.py:39-68``) for one image: returns ``y`` (float64, host)."""
    eng = engine or get_engine()
    torch = eng.torch
    g = np.ascontiguousarray(g_raw, dtype=np.float64)
    if g.ndim != 2: raise SolverError('preprocessing works on 2-D images')
    H, W = g.shape
    w1, r1 = gaussian_weights(sigma1)
    w2, r2 = gaussian_weights(sigma2)
    inf = float('inf')
    clip_abs = inf if np.isinf(offset_clip) else float(offset_clip * g.std())                  # preprocess.py:53
    if clip_abs < inf and not (g > clip_abs).any():
        raise SolverError('no pixel above offset_clip * std: the blend weights of preprocess.py:58-59 are 0 / 0 there')
    lower = float(g.mean()) if lower_clip_mean else -inf                                       # preprocess.py:61-62
    n = H * W
    d_g = eng._to_device('g_raw', g)
    d_y = eng._buf('d_pre_y', 8 * n)
    d_scratch = eng._buf('d_pre_scratch', 42 * n + 64)
    eng._check(eng.lib.sdsm_preprocess(eng.ctx, d_g.data_ptr(), H, W, w1.ctypes.data, r1, w2.ctypes.data, r2, float(sigma2), clip_abs, lower,
                                       d_y.data_ptr(), d_scratch.data_ptr(), eng.stream.cuda_stream))
    host = eng._buf('h_pre_y', 8 * n, pinned=True)
    with torch.cuda.stream(eng.stream):
        host[:8 * n].copy_(d_y[:8 * n], non_blocking=True)
    eng.stream.synchronize()
    eng.d2h_bytes += 8 * n
    return host[:8 * n].numpy().view(np.float64).reshape(H, W).copy()


class Preprocessing(object):
    """Stage-shaped wrapper (``name``, ``inputs``, ``outputs``, ``process``) with the reference's hyper-parameters
    ``preprocess/sigma1``, ``sigma2``, ``offset_clip``, ``lower_clip_mean`` and their defaults (``preprocess.py:41-47``)."""
    name, inputs, outputs = 'preprocess', ['g_raw'], ['y']

    def process(self, input_data, cfg, out=None, log_root_dir=None):
        return {'y': offset_image(input_data['g_raw'], cfg.get('sigma1', math.sqrt(2)), cfg.get('sigma2', 40),
                                  cfg.get('offset_clip', 3), cfg.get('lower_clip_mean', False))}

    def configure_ex(self, scale, radius, diameter):
        return {'sigma2': (scale, 1.0)}
