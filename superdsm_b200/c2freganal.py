"""The hot-path entry of the reference's coarse-to-fine region analysis (``superdsm/c2freganal.py:53-79``).

``C2F_RegionAnalysis`` decides every cluster split by comparing normalised energies of 6-parameter fits
(``smooth_amount = inf``, ``c2freganal.py:126``); this module mirrors the function that computes them, on top of the CUDA
``cvxprog``.  The split state machine itself (``c2freganal.py:193-287``), seeds and watershed stay host code of the
reference.
"""
import hashlib

import numpy as np

from .image import Image
from .objects import cvxprog


def _hash_mask(mask):
    """``c2freganal.py:53-55``."""
    mask = mask.astype(np.uint8)
    return hashlib.sha1(mask).digest()


def _get_cached_normalized_energy_computer(y, cluster):
    """``c2freganal.py:58-79``: ``compute_normalized_energy(obj, region, atoms_map, dsm_cfg) -> J(model) / n``.

    Same conventions as the reference: results are cached by the SHA-1 of the region mask; a region whose offset values
    are all positive or all negative has energy ``None``, and the division then raises ``TypeError`` exactly as the
    reference's does (its caller catches nothing: such regions do not occur behind ``c2freganal.py:215-230``).
    """
    cache = dict()
    cp_buffer = Image(model=y.model, mask=np.zeros(cluster.full_mask.shape, bool))

    def compute_normalized_energy(obj, region, atoms_map, dsm_cfg):
        cp_kwargs = dict(dsm_cfg)
        cp_kwargs.pop('smooth_mat_max_allocations', None)
        cp_region = obj.get_cvxprog_region(region, atoms_map, cp_kwargs.pop('background_margin'))
        cp_region_hash = _hash_mask(cp_region.mask)
        cache_hit = cache.get(cp_region_hash, None)
        if cache_hit is None:
            values = cp_region.model[cp_region.mask]
            if (values > 0).all() or (values < 0).all():
                energy = None
            else:
                cp_buffer.mask[cluster.full_mask] = cp_region.mask[cluster.mask]
                J, model, status = cvxprog(cp_buffer, smooth_mat_allocation_lock=None, **cp_kwargs)
                energy = J(model)
            cache_hit = energy / cp_region.mask.sum()
            cache[cp_region_hash] = cache_hit
        return cache_hit
    return compute_normalized_energy
