"""The split state machine of one cluster of the c2f stage as a GENERATOR (test infrastructure; my own statement of the loop
of ``superdsm/c2freganal.py:193-287``, not the reference's code): wherever the reference calls
``compute_normalized_energy(obj, region, atoms_map, dsm_cfg)`` this yields the region(s) instead and is sent the energies,
which is what lets ``superdsm_b200.c2f.run_level_synchronous`` batch all clusters of an image.

scikit-image is not installed, so the marker-based watershed is a small priority-flood implementation (4-connectivity).
"""
import heapq
import math

import numpy as np
import scipy.ndimage as ndi

from superdsm_b200.c2f import RegionRequest
from tests.compat.refshim import Image, Object

CROSS = ndi.generate_binary_structure(2, 1)


def watershed(image, markers, mask):
    """Marker-based watershed by priority flood: pixels are taken in order of (value, arrival) and get the label of the
    neighbour they were reached from."""
    out = np.where(mask, markers, 0).astype(np.int32)
    H, W = image.shape
    heap, tick = [], 0
    for r, c in zip(*np.nonzero(out)):
        heap.append((image[r, c], tick, r, c)); tick += 1
    heapq.heapify(heap)
    while heap:
        _, _, r, c = heapq.heappop(heap)
        for dr, dc in ((-1, 0), (1, 0), (0, -1), (0, 1)):
            rr, cc = r + dr, c + dc
            if 0 <= rr < H and 0 <= cc < W and mask[rr, cc] and out[rr, cc] == 0:
                out[rr, cc] = out[r, c]
                heapq.heappush(heap, (image[rr, cc], tick, rr, cc)); tick += 1
    return out


def next_seed(region, where, score):
    """The best local maximum of the region's image inside `where` (c2freganal.py:15-29, 4-connectivity)."""
    mask = region.mask & where
    peaks = (ndi.maximum_filter(region.model, footprint=CROSS) == region.model) & mask
    if not peaks.any(): return None
    lab, n = ndi.label(peaks)
    best, best_score = None, -np.inf
    for k in range(1, n + 1):
        s = score(lab == k)
        if s > best_score: best, best_score = lab == k, s
    return best if best_score > -np.inf else None


def cluster_task(y, clusters, cluster_label, y_mask, dsm_cfg, max_atom_norm_energy, min_atom_radius, min_norm_energy_improvement, log=None):
    """Generator.  Yields lists of RegionRequest, is sent lists of energies (float, None, or an exception instance); returns
    (cluster_label, atoms_map of the cluster crop, leaf energies)."""
    margin = dsm_cfg['background_margin']
    min_atom_size = math.pi * min_atom_radius ** 2
    cluster = y.get_region(clusters == cluster_label, shrink=True)
    masked = cluster.get_region(cluster.shrink_mask(y_mask))
    atoms_map = cluster.mask.astype(int)

    def request(obj):
        reg = obj.get_cvxprog_region(masked, atoms_map, margin)
        return RegionRequest(cluster.offset, reg.mask)

    def value(e):
        return None if (e is None or isinstance(e, Exception)) else e

    root = Object({1})
    root.seed = next_seed(masked, cluster.model > 0, lambda loc: cluster.model[loc].max())
    if root.seed is None: return cluster_label, atoms_map, []
    (e,) = yield [request(root)]
    if log is not None: log.append(('root', cluster_label, e))
    root.normalized_energy = value(e)
    if root.normalized_energy is None: return cluster_label, atoms_map, []
    leaves, queue = [], []
    (queue if root.normalized_energy > max_atom_norm_energy else leaves).append(root)
    seed_dist = ndi.distance_transform_edt(~root.seed)
    attempts = 0
    while queue and attempts < 40:
        attempts += 1
        c0 = queue.pop(0)
        c0_mask = c0.get_mask(atoms_map)
        if c0_mask.sum() < 2 * min_atom_size:
            leaves.append(c0); continue
        seed2 = next_seed(masked, (cluster.model > 0) & c0_mask & (seed_dist >= 1), lambda loc: seed_dist[loc].max())
        if seed2 is None:
            leaves.append(c0); continue
        seed_dist = np.minimum(seed_dist, ndi.distance_transform_edt(~seed2))
        sub = cluster.get_region(c0_mask)
        markers = np.zeros(sub.model.shape, int)
        markers[c0.seed] = 1
        markers[seed2] = 2
        ws = watershed(sub.model.max() - sub.model.clip(0, np.inf), markers, sub.mask)
        m1, m2 = ws == 1, ws == 2
        if m1.sum() < min_atom_size:
            c0.seed = seed2; queue.append(c0); continue
        if m2.sum() < min_atom_size:
            queue.append(c0); continue
        previous = atoms_map.copy()
        new_label = atoms_map.max() + 1
        atoms_map[m2] = new_label
        c1, c2 = Object(c0.footprint), Object({new_label})
        c1.seed, c2.seed = c0.seed, seed2
        e1, e2 = yield [request(c1), request(c2)]
        if log is not None: log.append(('split', cluster_label, e1, e2))
        c1.normalized_energy, c2.normalized_energy = value(e1), value(e2)
        if c1.normalized_energy is None or c2.normalized_energy is None:
            if c1.normalized_energy is None and c2.normalized_energy is not None: c0.seed = seed2
            atoms_map = previous
            queue.append(c0); continue
        improvement = 1 - max(c1.normalized_energy, c2.normalized_energy) / c0.normalized_energy
        if improvement < min_norm_energy_improvement:
            atoms_map = previous
            queue.append(c0)
        else:
            for c in (c1, c2):
                (queue if c.normalized_energy > max_atom_norm_energy else leaves).append(c)
    leaves += queue
    return cluster_label, atoms_map, [c.normalized_energy for c in leaves]
