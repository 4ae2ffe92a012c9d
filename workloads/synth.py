"""Synthetic fluorescence-nuclei workloads of the shapes named in BASELINE.json (the datasets are not available
offline): Gaussian-blob nuclei, multiplicative low-frequency illumination, additive noise, optional blur.

Produces what the hot path consumes: the offset image ``y`` (as ``preprocess.py:47-64`` of the reference defines it),
``y_mask``, an ``atoms`` label map that partitions the whole image, the atom adjacency, and candidate footprints
(atoms, cluster universes, sets of adjacent atoms as ``globalenergymin.py:292-307`` enumerates them).
The reference's c2f stage is not reproduced here: atoms are a nearest-seed partition of
each cluster, with nuclei optionally over-segmented into several atoms.
"""
import math

import numpy as np
import scipy.ndimage as ndi

#: name -> (shape, n_nuclei, (rmin, rmax), clustered fraction, blur, noise, illumination, AF scale or None)
WORKLOADS = {
    'bbbc039':  dict(shape=(520, 696),   n=100,   radius=(12, 20), clustered=0.2, blur=1.0, noise=0.02, illum=0.2, seed=39),
    'gowt1':    dict(shape=(1024, 1024), n=25,    radius=(45, 70), clustered=0.3, blur=6.0, noise=0.06, illum=0.3, seed=1),
    'nih3t3':   dict(shape=(1344, 1024), n=60,    radius=(35, 55), clustered=0.4, blur=2.0, noise=0.03, illum=0.75, seed=3),
    'tiled4k':  dict(shape=(4096, 4096), n=20000, radius=(10, 12), clustered=0.2, blur=1.0, noise=0.02, illum=0.2, seed=5),
    'toy':      dict(shape=(160, 200),   n=6,     radius=(14, 20), clustered=0.5, blur=1.0, noise=0.02, illum=0.1, seed=7),
}


def dsm_config(scale=None, **overrides):
    """``dsm_cfg`` as ``DSM_Config.process`` yields it (dsmcfg.py:6-21,82-97); ``scale`` applies the automation
    factors (``AF_scale``), None keeps the library defaults."""
    cfg = {
        'cachesize': 1, 'cachetest': None, 'sparsity_tol': 0, 'init': 'elliptical', 'smooth_amount': 10,
        'epsilon': 1.0, 'alpha': 0.5, 'scale': 1000, 'smooth_subsample': 20, 'gaussian_shape_multiplier': 2,
        'smooth_mat_dtype': 'float32', 'smooth_mat_max_allocations': np.inf, 'background_margin': 20, 'cp_timeout': 300,
    }
    if scale is not None:
        cfg['alpha'] = 0.0005 * scale ** 2
        cfg['smooth_amount'] = max(4, int(0.2 * scale))
        cfg['smooth_subsample'] = max(8, int(0.4 * scale))
        cfg['background_margin'] = max(8, int(0.4 * scale))
    cfg.update(overrides)
    return cfg


WORKLOAD_DSM = {
    'bbbc039': dict(scale=None),                       # BASELINE config 0: create_default_pipeline defaults
    'gowt1':   dict(scale=42.43, alpha=0.0005 * 42.43 ** 2),
    'nih3t3':  dict(scale=40, alpha=0.6),
    'tiled4k': dict(scale=10, alpha=3.3e-4 * 100),
    'toy':     dict(scale=10, alpha=3.3e-4 * 100),
}


def make_image(shape, n, radius, clustered, blur, noise, illum, seed):
    """Returns (g_raw in [0, 1], seed points, radii)."""
    rng = np.random.default_rng(seed)
    H, W = shape
    rmin, rmax = radius
    centres, radii = [], []
    cell = 2.6 * rmax
    gh, gw = max(1, int(H // cell)), max(1, int(W // cell))
    slots = rng.permutation(gh * gw)
    n_groups = min(len(slots), n)
    for slot in slots:
        if len(centres) >= n: break
        cy = (slot // gw + 0.5) * H / gh + rng.uniform(-0.15, 0.15) * cell
        cx = (slot % gw + 0.5) * W / gw + rng.uniform(-0.15, 0.15) * cell
        group = 1
        if rng.random() < clustered: group = int(rng.integers(2, 4))
        for k in range(group):
            if len(centres) >= n: break
            r = rng.uniform(rmin, rmax)
            ang = rng.uniform(0, 2 * math.pi)
            off = 0 if k == 0 else rng.uniform(1.2, 1.5) * r
            centres.append((cy + off * math.sin(ang), cx + off * math.cos(ang)))
            radii.append(r)
    img = np.zeros(shape)
    for (cy, cx), r in zip(centres, radii):
        ry, rx = r * rng.uniform(0.8, 1.0), r * rng.uniform(0.8, 1.0)
        th = rng.uniform(0, math.pi)
        y0, y1 = max(0, int(cy - 2 * r)), min(H, int(cy + 2 * r) + 1)
        x0, x1 = max(0, int(cx - 2 * r)), min(W, int(cx + 2 * r) + 1)
        if y0 >= y1 or x0 >= x1: continue
        yy, xx = np.mgrid[y0:y1, x0:x1]
        u = (yy - cy) * math.cos(th) + (xx - cx) * math.sin(th)
        v = -(yy - cy) * math.sin(th) + (xx - cx) * math.cos(th)
        d2 = (u / ry) ** 2 + (v / rx) ** 2
        img[y0:y1, x0:x1] = np.maximum(img[y0:y1, x0:x1], rng.uniform(0.6, 1.0) * np.exp(-0.5 * d2 ** 2))
    if blur > 0: img = ndi.gaussian_filter(img, blur)
    yy, xx = np.mgrid[0:H, 0:W]
    field = 1 + illum * np.sin(2 * math.pi * yy / (1.7 * H) + 0.3) * np.cos(2 * math.pi * xx / (1.3 * W) + 1.1)
    img = img * field + 0.05 * field
    img = img + rng.normal(0, noise, shape)
    img = np.round(np.clip(img, 0, None) * 40000).astype(np.uint16).astype(float)
    img = (img - img.min()) / max(img.max() - img.min(), 1)
    keep = [(cy, cx, r) for (cy, cx), r in zip(centres, radii) if 0 <= cy < H and 0 <= cx < W]
    return img, np.array([(cy, cx) for cy, cx, _ in keep]), np.array([r for _, _, r in keep])


def offset_image(g_raw, sigma1=math.sqrt(2), sigma2=40, offset_clip=3):
    """The offset image intensities y (reference ``preprocess.py:47-64``)."""
    offset = ndi.gaussian_filter(g_raw, sigma2)
    if np.isfinite(offset_clip):
        clip_abs = offset_clip * g_raw.std()
        clipped = ndi.gaussian_filter(g_raw.clip(0, clip_abs), sigma2)
        blend = (sigma2 - ndi.distance_transform_edt(~(g_raw > clip_abs))).clip(0, np.inf)
        blend = (blend / blend.max()) ** 2 if blend.max() > 0 else blend
        offset = (1 - blend) * clipped + blend * offset
    return ndi.gaussian_filter(g_raw, sigma1) - offset


def make_atoms(y, seeds, oversegment=1, rng=None, min_fg_pixels=30, max_irregularity=0.2, fast=False):
    """Partitions the image into clusters (nearest foreground component) and atoms (nearest seed inside the cluster).

    Returns (y_mask, atoms int32 HxW with labels 1..A, cluster label per atom, adjacency dict label -> set, clusters map,
    seed point of every atom in label order).
    """
    rng = rng or np.random.default_rng(0)
    H, W = y.shape
    fg = y > 0
    comp, ncomp = ndi.label(fg)
    index = np.arange(1, ncomp + 1)
    sizes = ndi.sum(fg, comp, index=index) if ncomp else np.zeros(0)
    # irregular (noise) components are masked out, as c2freganal.py:129-142 does with the perimeter/area ratio
    boundary = np.logical_xor(fg, ndi.binary_erosion(fg, ndi.generate_binary_structure(2, 1)))
    perimeter = ndi.sum(boundary, comp, index=index) if ncomp else np.zeros(0)
    y_mask = np.ones(y.shape, bool)
    bad = index[(sizes < min_fg_pixels) | (perimeter > max_irregularity * sizes)]
    if len(bad):
        lut = np.zeros(ncomp + 1, bool)
        lut[bad] = True
        y_mask[lut[comp]] = False
    markers = np.where(y_mask, comp, 0)
    if markers.max() == 0:
        return y_mask, np.zeros(y.shape, np.int32), np.zeros(0, int), {}, np.zeros(y.shape, np.int32), []
    idx = ndi.distance_transform_edt(markers == 0, return_distances=False, return_indices=True)
    clusters = markers[idx[0], idx[1]]
    # seeds: given nuclei centres (those lying on foreground) plus, per cluster without any, its brightest pixel
    pts = [(int(round(sy)), int(round(sx))) for sy, sx in seeds]
    pts = [(r, c) for r, c in pts if 0 <= r < H and 0 <= c < W and markers[r, c] > 0]
    extra = []
    for (r, c) in pts:
        for _ in range(oversegment - 1):
            rr, cc = r + int(rng.integers(-6, 7)), c + int(rng.integers(-6, 7))
            if 0 <= rr < H and 0 <= cc < W and markers[rr, cc] == markers[r, c]: extra.append((rr, cc))
    pts = pts + extra
    have = set(int(markers[r, c]) for r, c in pts)
    if fast:
        missing = [int(lab) for lab in np.unique(markers) if lab != 0 and int(lab) not in have]
        if missing:
            for pos in ndi.maximum_position(y, markers, index=missing): pts.append((int(pos[0]), int(pos[1])))
    else:
        for lab in np.unique(markers):
            if lab == 0 or int(lab) in have: continue
            pos = ndi.maximum_position(y, markers, lab)
            pts.append((int(pos[0]), int(pos[1])))
    pts = sorted(set(pts))
    seed_map = np.zeros(y.shape, np.int32)
    for k, (r, c) in enumerate(pts, start=1): seed_map[r, c] = k
    # nearest seed of the same cluster: a Voronoi partition per cluster
    atoms = np.zeros(y.shape, np.int32)
    seed_cluster = np.array([clusters[r, c] for r, c in pts])
    if fast:
        # the same partition computed on the bounding box of every cluster (large images: the loop below works on full frames)
        by_cluster = {}
        for k, lab in enumerate(seed_cluster): by_cluster.setdefault(int(lab), []).append(k)
        for lab, sl in enumerate(ndi.find_objects(clusters), start=1):
            if sl is None or lab not in by_cluster: continue
            sel = by_cluster[lab]
            region = clusters[sl] == lab
            if len(sel) == 1:
                atoms[sl][region] = sel[0] + 1
                continue
            sub = np.ones(region.shape, bool)
            for k in sel: sub[pts[k][0] - sl[0].start, pts[k][1] - sl[1].start] = False
            idx2 = ndi.distance_transform_edt(sub, return_distances=False, return_indices=True)
            atoms[sl][region] = seed_map[sl][idx2[0][region], idx2[1][region]]
    for lab in (np.unique(clusters) if not fast else []):
        sel = np.nonzero(seed_cluster == lab)[0]
        region = clusters == lab
        if len(sel) == 1:
            atoms[region] = sel[0] + 1
            continue
        sub = np.ones(y.shape, bool)
        for k in sel: sub[pts[k]] = False
        idx2 = ndi.distance_transform_edt(sub, return_distances=False, return_indices=True)
        atoms[region] = seed_map[idx2[0][region], idx2[1][region]]
    # relabel 1..A in order of first appearance by label value
    labels = np.unique(atoms)
    labels = labels[labels > 0]
    lut = np.zeros(atoms.max() + 1, np.int32)
    lut[labels] = np.arange(1, len(labels) + 1)
    atoms = lut[atoms]
    atom_cluster = np.zeros(len(labels) + 1, int)
    atom_cluster[atoms.reshape(-1)] = clusters.reshape(-1)
    adjacency = {int(a): set() for a in range(1, len(labels) + 1)}
    for a, b in ((atoms[:-1, :], atoms[1:, :]), (atoms[:, :-1], atoms[:, 1:])):
        diff = (a != b) & (a > 0) & (b > 0)
        pairs = np.unique(np.stack([a[diff], b[diff]], axis=1), axis=0) if diff.any() else np.zeros((0, 2), int)
        for u, v in pairs:
            if atom_cluster[u] == atom_cluster[v]:
                adjacency[int(u)].add(int(v))
                adjacency[int(v)].add(int(u))
    return y_mask, atoms, atom_cluster[1:], adjacency, clusters.astype(np.int32), [pts[int(k) - 1] for k in labels]


def enumerate_candidates(atom_cluster, adjacency, max_set_size=3, max_candidates=None):
    """Footprints in the order the reference's GEM stage solves them: atoms, cluster universes, then generations of
    sets of adjacent atoms (``globalenergymin.py:183-307``), without pruning."""
    n_atoms = len(atom_cluster)
    atoms = [frozenset([a]) for a in range(1, n_atoms + 1)]
    universes = []
    for lab in np.unique(atom_cluster):
        members = frozenset(int(a) + 1 for a in np.nonzero(atom_cluster == lab)[0])
        if len(members) > 1: universes.append(members)
    seen = set(atoms) | set(universes)
    generations = []
    prev = atoms
    for size in range(2, max_set_size + 1):
        nxt = []
        for fp in prev:
            adj = set()
            for a in fp: adj |= adjacency[a] - fp
            for b in sorted(adj):
                new = frozenset(fp | {b})
                if new not in seen:
                    seen.add(new)
                    nxt.append(new)
        if not nxt: break
        generations.append(nxt)
        prev = nxt
    out = atoms + universes + [fp for gen in generations for fp in gen]
    if max_candidates is not None: out = out[:max_candidates]
    return out


def make_workload(name, oversegment=2, max_set_size=3, max_candidates=None, seed=None, shape=None, n=None):
    """Returns dict(y, y_mask, atoms, footprints, dsm_cfg, name)."""
    spec = dict(WORKLOADS[name])
    if seed is not None: spec['seed'] = seed
    if shape is not None: spec['shape'] = tuple(shape)
    if n is not None: spec['n'] = n
    g_raw, seeds, radii = make_image(**spec)
    dsm_kwargs = dict(WORKLOAD_DSM[name])
    scale = dsm_kwargs.pop('scale')
    dsm_cfg = dsm_config(scale, **dsm_kwargs)
    y = offset_image(g_raw, sigma2=(40 if scale is None else scale))
    y_mask, atoms, atom_cluster, adjacency, clusters, atom_seeds = make_atoms(y, seeds, oversegment, np.random.default_rng(spec['seed'] + 1))
    footprints = enumerate_candidates(atom_cluster, adjacency, max_set_size, max_candidates)
    return dict(name=name, g_raw=g_raw, y=y, y_mask=y_mask, atoms=atoms, footprints=footprints, dsm_cfg=dsm_cfg,
                adjacency=adjacency, atom_cluster=atom_cluster, clusters=clusters, atom_seeds=atom_seeds)
