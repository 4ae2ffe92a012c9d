"""Host driver of the CUDA solver: owns the device buffers (torch tensors), prepares the candidate descriptors and
calls through the C ABI (``include/superdsm_b200.h``).  One ``Engine`` per device; one process per GPU.

PyTorch is used for device memory, pinned staging buffers and streams only.  There is no CPU path.
"""
import ctypes
import time

import numpy as np
import scipy.ndimage as ndi

from . import _lib


class SolverError(RuntimeError):
    pass


def gaussian_psf32(smooth_amount, gaussian_shape_multiplier):
    """The float32 Gaussian kernel of the reference (``dsm.py:137-142,226``) and its radius R.

    Built exactly as the reference builds it (``scipy.ndimage.gaussian_filter`` of a centred delta), so the values are
    bit-identical by construction; the device receives the quarter ``psf[R:, R:]`` of the symmetric box.
    """
    size = round(1 + smooth_amount * 4 * gaussian_shape_multiplier)
    assert size % 2 == 1, size                              # dsm.py:147
    inp = np.zeros((size, size))
    inp[size // 2, size // 2] = 1
    psf = ndi.gaussian_filter(inp, smooth_amount).astype(np.float32)
    R = size // 2
    if not (np.array_equal(psf, psf[::-1, :]) and np.array_equal(psf, psf[:, ::-1])):
        raise SolverError('asymmetric Gaussian kernel: not supported')
    return np.ascontiguousarray(psf[R:, R:]), R


def lpt_shards(costs, n_shards):
    """Longest-processing-time assignment of candidates to shards (ranks); returns a list of index arrays."""
    order = np.argsort(-np.asarray(costs, dtype=float), kind='stable')
    loads = np.zeros(n_shards)
    shards = [[] for _ in range(n_shards)]
    for k in order:
        r = int(np.argmin(loads))
        shards[r].append(int(k))
        loads[r] += costs[k]
    return [np.array(sorted(s), dtype=np.int64) for s in shards]


def prepare_image(y_model, y_mask, atoms, background_margin):
    """Host-side, once per image: the candidate-independent part of ``Object.get_cvxprog_region`` (objects.py:126-127).

    A pixel is usable iff ``y_mask`` and ``EDT(y <= 0) <= background_margin``; unusable pixels get atom label 0, so that a
    candidate's region is simply {pixels whose label is in its footprint}.  Returns the masked atoms map (int32), the
    number of usable pixels per label and the bounding box (r0, r1, c0, c1; end-exclusive) of each label's usable pixels.
    """
    atoms = np.asarray(atoms)
    assert y_model.ndim == 2 and atoms.shape == y_model.shape and y_mask.shape == y_model.shape
    usable = np.logical_and(y_mask, ndi.distance_transform_edt(y_model <= 0) <= background_margin)
    atoms_v = np.where(usable, atoms, 0).astype(np.int32)
    n_labels = int(atoms.max()) if atoms.size else 0
    count = np.bincount(atoms_v.reshape(-1), minlength=n_labels + 1).astype(np.int64)
    count[0] = 0
    box = np.zeros((n_labels + 1, 4), dtype=np.int64)
    for lab, sl in enumerate(ndi.find_objects(atoms_v, max_label=n_labels), start=1):
        if sl is not None: box[lab] = (sl[0].start, sl[0].stop, sl[1].start, sl[1].stop)
    return atoms_v, count, box


def describe_candidates(atom_count, atom_box, footprints):
    """Candidate descriptors (``sdsm_cand``: bounding box, pixel count, footprint slice) and the flat footprint array."""
    n = len(footprints)
    lens = np.fromiter((len(fp) for fp in footprints), dtype=np.int64, count=n)
    if n == 0 or (lens < 1).any(): raise SolverError('empty batch or empty footprint')
    flat = np.fromiter((lab for fp in footprints for lab in sorted(fp)), dtype=np.int64, count=int(lens.sum()))
    if flat.min() < 1 or flat.max() >= len(atom_count):
        raise SolverError('footprint label outside the atoms map')
    offs = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int64)
    cnt = atom_count[flat]
    box = atom_box[flat]
    big = np.iinfo(np.int64).max
    present = cnt > 0
    r0 = np.minimum.reduceat(np.where(present, box[:, 0], big), offs)
    r1 = np.maximum.reduceat(np.where(present, box[:, 1], 0), offs)
    c0 = np.minimum.reduceat(np.where(present, box[:, 2], big), offs)
    c1 = np.maximum.reduceat(np.where(present, box[:, 3], 0), offs)
    npix = np.add.reduceat(cnt, offs)
    if (npix < 1).any():
        raise SolverError(f'candidate {int(np.argmax(npix < 1))} has no usable pixel')
    cands = np.zeros(n, dtype=_lib.CAND_DTYPE)
    cands['r0'], cands['c0'], cands['h'], cands['w'] = r0, c0, r1 - r0, c1 - c0
    cands['fp_off'], cands['fp_len'], cands['n'] = offs, lens, npix
    return cands, flat.astype(np.int32)


class BatchResult(object):
    """Raw per-candidate results of one batch (host numpy arrays)."""

    def __init__(self, cands, energy, status, info, params, fgbox, on_boundary, fgmask, elapsed):
        self.cands, self.energy, self.status, self.info = cands, energy, status, info
        self.params_flat, self.fgbox, self.on_boundary, self.fgmask = params, fgbox, on_boundary, fgmask
        self.elapsed = elapsed

    def __len__(self): return len(self.cands)

    def m(self, k): return int(self.info[k, 0])

    def params(self, k):
        off = int(self.cands['par_off'][k])
        return self.params_flat[off:off + 6 + self.m(k)].copy()

    def foreground(self, k):
        """(fg_offset, fg_fragment) as ``objects.extract_foreground_fragment`` returns them (objects.py:148-174)."""
        rmin, rmax, cmin, cmax = (int(v) for v in self.fgbox[k])
        if rmax < 0: return np.zeros(2, int), np.zeros((1, 1), bool)
        cd = self.cands[k]
        h, w, off = int(cd['h']), int(cd['w']), int(cd['fg_off'])
        box = self.fgmask[off:off + h * w].reshape(h, w)
        return np.array([int(cd['r0']) + rmin, int(cd['c0']) + cmin]), box[rmin:rmax + 1, cmin:cmax + 1].astype(bool)


class DeviceImage(object):
    """One image resident on the device: y, its mask, the atoms map, the masked atoms map the kernels read
    (``sdsm_prepare_image``) and, on the host, the per-atom usable-pixel counts and bounding boxes."""

    def __init__(self, shape):
        self.shape = tuple(shape)
        self.d_y = self.d_mask = self.d_atoms_in = self.d_atoms = self.d_count = self.d_box = self.d_scratch = None
        self.atom_count = self.atom_box = None
        self.n_labels = 0


class Engine(object):
    """One device context of the CUDA solver."""

    def __init__(self, device=0):
        import torch
        if not torch.cuda.is_available():
            raise SolverError('superdsm_b200 needs a CUDA device (B200); there is no CPU fallback')
        self.torch = torch
        self.lib = _lib.load()
        self.device = torch.device('cuda', device)
        self.device_index = device
        self.ctx = ctypes.c_void_p()
        rc = self.lib.sdsm_create(ctypes.byref(self.ctx), device)
        self._check(rc)
        self.stream = torch.cuda.Stream(device=self.device)
        self._params_key = None
        self._bufs = {}
        self._staged = {}
        self._last = None
        self._image = None             # the DeviceImage of set_image()
        self._bound = []               # the images the context currently solves on
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    def close(self):
        if self.ctx:
            self.torch.cuda.synchronize(self.device)
            self.lib.sdsm_destroy(self.ctx)
            self.ctx = ctypes.c_void_p()

    def __del__(self):
        try: self.close()
        except Exception: pass

    def _check(self, rc):
        if rc != _lib.SDSM_OK:
            msg = self.lib.sdsm_last_error(self.ctx)
            raise SolverError(f'superdsm_b200 error {rc}: {msg.decode() if msg else ""}')

    @property
    def launch_count(self):
        return int(self.lib.sdsm_launch_count(self.ctx))

    STAT_NAMES = ('q_head', 'q_tail', 'done', 'abort', 'cyc_eval', 'cyc_hrows', 'cyc_ctrl', 'cyc_idle', 'n_eval', 'n_hrows',
                  'n_ctrl', 'cyc_total', 't0', 'cyc_widle', 'cyc_wtotal', 'cyc_wfin')

    def solver_stats(self, ws=None, plan=None):
        """Counters of the persistent solver kernel of the last batch (synchronises): see ``sdsm_solver_stats``."""
        if ws is None: plan, ws = self._last[0], self._last[2]
        st = np.zeros(32, dtype=np.uint64)
        self._check(self.lib.sdsm_solver_stats(self.ctx, ctypes.byref(plan), ws.data_ptr(), st.ctypes.data, self.stream.cuda_stream))
        out = {k: int(st[i]) for i, k in enumerate(self.STAT_NAMES)}
        for j, kind in enumerate(('poly_f', 'poly_h', 'xi_f', 'xi_h')):
            out['cyc_' + kind], out['n_' + kind] = int(st[16 + j]), int(st[20 + j])
        out['q_head_hi'], out['q_tail_hi'], out['n_promoted'] = int(st[24]), int(st[25]), int(st[26])
        out['cyc_ctrl_big'], out['n_ctrl_big'] = int(st[27]), int(st[28])
        out['cyc_chol_big'], out['n_chol_big'], out['cyc_solve_big'] = int(st[29]), int(st[30]), int(st[31])
        return out

    def finish_times_us(self, ws=None, n=None, plan=None):
        """Per-candidate finish time of the last batch, microseconds after the solver kernel started (synchronises)."""
        if ws is None: plan, ws, n = self._last[0], self._last[2], len(self._last[3])
        out = np.zeros(2 * n, dtype=np.int32)
        self._check(self.lib.sdsm_finish_times(self.ctx, ctypes.byref(plan), ws.data_ptr(), out.ctypes.data, self.stream.cuda_stream))
        self.control_times_us = out[n:]
        return out[:n]

    def eval_counts(self, ws=None, n=None, plan=None):
        """Per candidate of the last batch (n x 4, synchronises): trials decided without evaluating F, trial points evaluated
        ahead and never examined, accepted trials evaluated together with their Hessian, Hessians evaluated in vain
        (``sdsm_eval_counts``; ``info[:, 6:8]`` count the evaluations the reference makes)."""
        if ws is None: plan, ws, n = self._last[0], self._last[2], len(self._last[3])
        out = np.zeros((n, 4), dtype=np.int32)
        self._check(self.lib.sdsm_eval_counts(self.ctx, ctypes.byref(plan), ws.data_ptr(), out.ctypes.data, self.stream.cuda_stream))
        return out

    # -- buffers -------------------------------------------------------------------------------------------------

    def _buf(self, name, nbytes, pinned=False):
        """Grow-only byte buffers, device or pinned host."""
        torch = self.torch
        cur = self._bufs.get(name)
        if cur is None or cur.numel() < nbytes:
            cap = max(int(nbytes * 1.25), 1024)
            if pinned: cur = torch.empty(cap, dtype=torch.uint8, pin_memory=True)
            else: cur = torch.empty(cap, dtype=torch.uint8, device=self.device)
            self._bufs[name] = cur
        return cur

    def _to_device(self, name, array, dev=None):
        """numpy -> pinned staging -> device tensor `dev` (or the grow-only buffer 'd_' + name), on the engine's stream.
        The staging buffer of a name is reused only after the copy that last read it has finished (an event per name: no
        stream-wide synchronisation)."""
        torch = self.torch
        array = np.ascontiguousarray(array)
        nbytes = array.nbytes
        ev = self._staged.get(name)
        if ev is not None: ev.synchronize()
        host = self._buf('h_' + name, nbytes, pinned=True)
        host[:nbytes].numpy()[:] = array.reshape(-1).view(np.uint8)
        if dev is None: dev = self._buf('d_' + name, nbytes)
        with torch.cuda.stream(self.stream):
            dev[:nbytes].copy_(host[:nbytes], non_blocking=True)
            if ev is None: ev = self._staged[name] = torch.cuda.Event()
            ev.record(self.stream)
        self.h2d_bytes += nbytes
        return dev

    # -- images --------------------------------------------------------------------------------------------------

    def upload_image(self, y_model, y_mask, atoms, background_margin, image=None):
        """Uploads the offset image, its mask and the atoms map into a :class:`DeviceImage` (a new one, or `image` to reuse
        its buffers); the usable pixels (``objects.py:126-127``: mask and ``EDT(y <= 0) <= background_margin``), the masked
        atoms map and the per-atom pixel counts / bounding boxes are computed on the device (``sdsm_prepare_image``;
        :func:`prepare_image` is the numpy statement of the same thing)."""
        torch = self.torch
        y_model = np.ascontiguousarray(y_model, dtype=np.float64)
        assert y_model.ndim == 2
        H, W = y_model.shape
        if image is None or image.shape != (H, W):
            image = DeviceImage((H, W))
            dev = lambda n: torch.empty(n, dtype=torch.uint8, device=self.device)
            image.d_y, image.d_mask, image.d_atoms_in, image.d_atoms, image.d_scratch = dev(8 * H * W), dev(H * W), dev(4 * H * W), dev(4 * H * W), dev(2 * H * W)
        self._to_device('y', y_model, image.d_y)
        return self.update_labels(image, y_mask, atoms, background_margin)

    def update_labels(self, image, y_mask, atoms, background_margin):
        """New mask / atoms map for an image whose y stays resident (the rounds of the level-synchronous c2f scheduler):
        re-runs ``sdsm_prepare_image``."""
        torch = self.torch
        H, W = image.shape
        atoms = np.ascontiguousarray(atoms, dtype=np.int32)
        y_mask = np.ascontiguousarray(y_mask, dtype=np.bool_)
        assert atoms.shape == (H, W) and y_mask.shape == (H, W)
        n_labels = int(atoms.max()) if atoms.size else 0
        self._to_device('ymask', y_mask.view(np.uint8), image.d_mask)
        self._to_device('atoms_in', atoms, image.d_atoms_in)
        if image.d_count is None or image.d_count.numel() < 8 * (n_labels + 1):
            image.d_count = torch.empty(8 * (n_labels + 1) * 2, dtype=torch.uint8, device=self.device)
            image.d_box = torch.empty(16 * (n_labels + 1) * 2, dtype=torch.uint8, device=self.device)
        margin = float(background_margin)
        self._check(self.lib.sdsm_prepare_image(self.ctx, image.d_y.data_ptr(), image.d_mask.data_ptr(), image.d_atoms_in.data_ptr(), H, W,
                                                margin if np.isfinite(margin) else 1e308 * 10, n_labels, image.d_atoms.data_ptr(),
                                                image.d_count.data_ptr(), image.d_box.data_ptr(), image.d_scratch.data_ptr(), self.stream.cuda_stream))
        nb_c, nb_b = 8 * (n_labels + 1), 16 * (n_labels + 1)
        h_c, h_b = self._buf('h_atom_count', nb_c, pinned=True), self._buf('h_atom_box', nb_b, pinned=True)
        with torch.cuda.stream(self.stream):
            h_c[:nb_c].copy_(image.d_count[:nb_c], non_blocking=True)
            h_b[:nb_b].copy_(image.d_box[:nb_b], non_blocking=True)
        self.stream.synchronize()                      # the descriptors of the candidates need the counts and boxes
        self.d2h_bytes += nb_c + nb_b
        image.n_labels = n_labels
        image.atom_count = h_c[:nb_c].numpy().view(np.int64).copy()
        box = h_b[:nb_b].numpy().view(np.int32).reshape(-1, 4).astype(np.int64)
        box[image.atom_count == 0] = 0
        image.atom_box = box
        return image

    def bind_images(self, images):
        """The images the following batches draw their candidates from (``sdsm_set_images``)."""
        images = list(images)
        if not 1 <= len(images) <= _lib.MAX_IMAGES: raise SolverError(f'1 .. {_lib.MAX_IMAGES} images per batch')
        arr = (_lib.ImageDesc * len(images))()
        for k, im in enumerate(images):
            arr[k].d_y, arr[k].d_atoms, arr[k].H, arr[k].W = im.d_y.data_ptr(), im.d_atoms.data_ptr(), im.shape[0], im.shape[1]
        self._check(self.lib.sdsm_set_images(self.ctx, arr, len(images)))
        self._bound = images
        self._last = None

    def set_image(self, y_model, y_mask, atoms, background_margin):
        """Single-image convenience: uploads and binds one image (the engine keeps and reuses its device buffers)."""
        self._image = self.upload_image(y_model, y_mask, atoms, background_margin, image=self._image)
        self.shape, self.atom_count, self.atom_box = self._image.shape, self._image.atom_count, self._image.atom_box
        self.bind_images([self._image])

    def masked_atoms(self, image=None):
        """The masked atoms map of an image (device -> host; diagnostics and tests)."""
        image = image or self._bound[0]
        H, W = image.shape
        self.stream.synchronize()
        return image.d_atoms[:4 * H * W].cpu().numpy().view(np.int32).reshape(H, W)

    def set_params(self, dsm_cfg, skip_single_fg=True):
        """``dsm_cfg`` is the dictionary of ``dsmcfg.py:82-89``.  ``skip_single_fg``: the rule of ``_compute_object``
        (objects.py:184-191); ``cvxprog`` itself (the c2f seam) fits such regions, so it passes False."""
        cfg = dict(dsm_cfg)
        if cfg.get('smooth_mat_dtype', 'float32') != 'float32':
            raise SolverError('only dsm/smooth_mat_dtype = float32 (the default) is implemented')
        if callable(cfg.get('init')):
            raise SolverError('callable dsm/init is not implemented on the device')
        key = tuple(sorted((k, str(v)) for k, v in cfg.items() if k in
                           ('alpha', 'epsilon', 'scale', 'sparsity_tol', 'hessian_sparsity_tol', 'smooth_amount', 'smooth_subsample',
                            'gaussian_shape_multiplier', 'init'))) + (bool(skip_single_fg),)
        if key == self._params_key: return
        p = _lib.Params()
        p.alpha, p.epsilon, p.scale = float(cfg['alpha']), float(cfg['epsilon']), float(cfg['scale'])
        p.sparsity_tol = float(cfg.get('sparsity_tol', 0))
        p.hessian_sparsity_tol = float(cfg.get('hessian_sparsity_tol', 0))
        p.subsample = int(cfg['smooth_subsample']) if np.isfinite(cfg['smooth_subsample']) else 16
        p.init_elliptical = 1 if cfg.get('init', 'elliptical') == 'elliptical' else 0
        p.maxiters, p.abstol, p.reltol, p.feastol = 100, 1e-7, 1e-6, 1e-7      # cvxopt defaults (never changed by the reference)
        p.skip_single_fg = 1 if skip_single_fg else 0
        self._psf = None
        if cfg['smooth_amount'] < np.inf:
            self._psf, R = gaussian_psf32(cfg['smooth_amount'], cfg['gaussian_shape_multiplier'])
            p.has_smooth, p.radius = 1, R
            p.h_psf = self._psf.ctypes.data
        else:
            p.has_smooth, p.radius, p.h_psf = 0, 0, None
        self._check(self.lib.sdsm_set_params(self.ctx, ctypes.byref(p)))
        self._params_key = key
        self._last = None

    # -- per batch -----------------------------------------------------------------------------------------------

    def describe(self, footprints, image=None):
        image = image or self._bound[0]
        return describe_candidates(image.atom_count, image.atom_box, footprints)

    def describe_multi(self, footprints_per_image):
        """Descriptors of a batch over the bound images: ``footprints_per_image[i]`` are the candidates of image i."""
        if len(footprints_per_image) != len(self._bound): raise SolverError('one footprint list per bound image')
        parts, flats, off = [], [], 0
        for i, (image, fps) in enumerate(zip(self._bound, footprints_per_image)):
            if len(fps) == 0: continue
            cands, flat = describe_candidates(image.atom_count, image.atom_box, fps)
            cands['img'] = i
            cands['fp_off'] += off
            off += len(flat)
            parts.append(cands); flats.append(flat)
        if not parts: raise SolverError('empty batch')
        return np.concatenate(parts), np.concatenate(flats)

    def _plan(self, cands):
        plan = _lib.Plan()
        self._check(self.lib.sdsm_plan_batch(self.ctx, cands.ctypes.data, len(cands), ctypes.byref(plan)))
        # most expensive first: the persistent solver pulls candidates from a queue
        cost = cands['n'].astype(np.float64) * (1.0 + 0.02 * cands['cell_cap'])
        return plan, np.argsort(-cost, kind='stable').astype(np.int32)

    @staticmethod
    def _out_sizes(n, plan):
        return dict(energy=8 * n, status=4 * n, info=32 * n, params=8 * plan.params_len, fgbox=16 * n, onb=n, fgmask=plan.fgmask_len, solver=16)

    def solve(self, footprints, keep_device_state=False, borrow=False):
        """Fits every candidate of the batch (candidates of the first bound image); returns a :class:`BatchResult`.
        ``borrow=True``: the foreground masks of the result are a view of the engine's pinned staging buffer (valid until
        the next ``solve``) instead of a copy."""
        cands, flat = self.describe(footprints)
        return self._solve_described(cands, flat, borrow)

    def solve_multi(self, footprints_per_image, borrow=False):
        """One batch over all bound images (``bind_images``); ``result.cands['img']`` tells the candidates apart."""
        cands, flat = self.describe_multi(footprints_per_image)
        return self._solve_described(cands, flat, borrow)

    def _solve_described(self, cands, flat, borrow):
        torch, lib = self.torch, self.lib
        t0 = time.perf_counter()
        n = len(cands)
        plan, order = self._plan(cands)
        d_cands = self._to_device('cands', cands)
        d_fp = self._to_device('fp', flat)
        d_order = self._to_device('order', order)
        ws = self._buf('d_ws', plan.workspace_bytes)
        sizes = self._out_sizes(n, plan)
        dev = {k: self._buf('d_out_' + k, v) for k, v in sizes.items()}
        out = _lib.Out(dev['energy'].data_ptr(), dev['status'].data_ptr(), dev['info'].data_ptr(), dev['params'].data_ptr(),
                       dev['fgbox'].data_ptr(), dev['onb'].data_ptr(), dev['fgmask'].data_ptr(), dev['solver'].data_ptr())
        self._check(lib.sdsm_solve_batch(self.ctx, ctypes.byref(plan), d_cands.data_ptr(), n, d_fp.data_ptr(), d_order.data_ptr(),
                                         ws.data_ptr(), ws.numel(), ctypes.byref(out), self.stream.cuda_stream))
        host = {}
        with torch.cuda.stream(self.stream):
            for k, v in sizes.items():
                h = self._buf('h_out_' + k, v, pinned=True)
                h[:v].copy_(dev[k][:v], non_blocking=True)
                host[k] = h
                self.d2h_bytes += v
        self.stream.synchronize()
        done, abort = (int(v) for v in host['solver'][:16].numpy().view(np.uint64))
        if abort or done != n:
            raise SolverError(f'persistent solver did not finish: {done} of {n} candidates done, abort={abort}')
        res = BatchResult(cands,
                          host['energy'][:sizes['energy']].numpy().view(np.float64).copy(),
                          host['status'][:sizes['status']].numpy().view(np.int32).copy(),
                          host['info'][:sizes['info']].numpy().view(np.int32).reshape(n, 8).copy(),
                          host['params'][:sizes['params']].numpy().view(np.float64).copy(),
                          host['fgbox'][:sizes['fgbox']].numpy().view(np.int32).reshape(n, 4).copy(),
                          host['onb'][:n].numpy().copy(),
                          host['fgmask'][:sizes['fgmask']].numpy() if borrow else host['fgmask'][:sizes['fgmask']].numpy().copy(),
                          time.perf_counter() - t0)
        self._last = (plan, d_cands, ws, cands)
        return res

    # -- device-resident variant used by bench.py: inputs already in HBM, outputs stay there -------------------------

    def prepare(self, footprints, multi=False):
        """Uploads a batch and returns a handle for :meth:`run_prepared` (no host work inside the timed region).  The handle owns
        its workspace and carries its plan: it stays valid whatever is planned or solved afterwards (on the same images)."""
        cands, flat = self.describe_multi(footprints) if multi else self.describe(footprints)
        n = len(cands)
        plan, order = self._plan(cands)
        torch = self.torch
        to_dev = lambda a: torch.from_numpy(np.ascontiguousarray(a).reshape(-1).view(np.uint8).copy()).to(self.device)
        h = dict(plan=plan, n=n, cands=cands, d_cands=to_dev(cands), d_fp=to_dev(flat), d_order=to_dev(order))
        h['ws'] = torch.empty(plan.workspace_bytes, dtype=torch.uint8, device=self.device)
        sizes = self._out_sizes(n, plan)
        h['out'] = {k: torch.empty(max(v, 8), dtype=torch.uint8, device=self.device) for k, v in sizes.items()}
        o = h['out']
        h['out_struct'] = _lib.Out(o['energy'].data_ptr(), o['status'].data_ptr(), o['info'].data_ptr(), o['params'].data_ptr(),
                                   o['fgbox'].data_ptr(), o['onb'].data_ptr(), o['fgmask'].data_ptr(), o['solver'].data_ptr())
        torch.cuda.synchronize(self.device)
        return h

    def run_prepared(self, h, stream=None):
        stream = stream if stream is not None else self.torch.cuda.current_stream(self.device)
        self._check(self.lib.sdsm_solve_batch(self.ctx, ctypes.byref(h['plan']), h['d_cands'].data_ptr(), h['n'], h['d_fp'].data_ptr(),
                                              h['d_order'].data_ptr(), h['ws'].data_ptr(), h['ws'].numel(),
                                              ctypes.byref(h['out_struct']), stream.cuda_stream))
        self._last = (h['plan'], h['d_cands'], h['ws'], h['cands'])

    # -- diagnostics (parity tests) ----------------------------------------------------------------------------------

    def eval_candidate(self, k, params, hessian=True):
        """psi, grad psi, Hessian of candidate ``k`` of the last batch at ``params`` (device evaluation).  ``hessian=False``
        runs the value + gradient evaluation the line searches use and returns ``None`` for the Hessian."""
        torch = self.torch
        plan, d_cands, ws, cands = self._last
        params = np.ascontiguousarray(params, dtype=np.float64)
        N = len(params)
        d_x = torch.from_numpy(params).to(self.device)
        d_val = torch.zeros(1, dtype=torch.float64, device=self.device)
        d_g = torch.zeros(N, dtype=torch.float64, device=self.device)
        d_H = torch.zeros(N * N, dtype=torch.float64, device=self.device)
        torch.cuda.synchronize(self.device)
        self._check(self.lib.sdsm_eval_candidate(self.ctx, ctypes.byref(plan), d_cands.data_ptr(), k, d_x.data_ptr(), d_val.data_ptr(),
                                                 d_g.data_ptr(), d_H.data_ptr() if hessian else None, ws.data_ptr(), self.stream.cuda_stream))
        self.stream.synchronize()
        return float(d_val.item()), d_g.cpu().numpy(), d_H.cpu().numpy().reshape(N, N) if hessian else None

    def trace_solve(self, k, x0, use_xi=True, maxiters=100):
        """One interior-point solve of candidate ``k`` of the last batch from ``x0``; returns (status, iterations, x, trace)."""
        torch = self.torch
        plan, d_cands, ws, cands = self._last
        x0 = np.ascontiguousarray(x0, dtype=np.float64)
        d_x0 = torch.from_numpy(x0).to(self.device)
        d_tr = torch.zeros((maxiters + 1) * 8, dtype=torch.float64, device=self.device)
        d_x = torch.zeros(len(x0), dtype=torch.float64, device=self.device)
        d_si = torch.zeros(2, dtype=torch.int32, device=self.device)
        torch.cuda.synchronize(self.device)
        self._check(self.lib.sdsm_trace_solve(self.ctx, ctypes.byref(plan), d_cands.data_ptr(), k, 1 if use_xi else 0, d_x0.data_ptr(),
                                              d_tr.data_ptr(), d_x.data_ptr(), d_si.data_ptr(), ws.data_ptr(), self.stream.cuda_stream))
        self.stream.synchronize()
        si = d_si.cpu().numpy()
        n_it = int(si[1])
        return int(si[0]), n_it, d_x.cpu().numpy(), d_tr.cpu().numpy().reshape(-1, 8)[:n_it + 1]

    def sweep_eval_ms(self, h, want_h=False, rows=False, repeats=5):
        """Average milliseconds of one sweep of the evaluation tasks (or Hessian-row tasks) over every candidate of the
        prepared batch ``h`` at its solution (after :meth:`run_prepared`); CUDA events on the launching stream."""
        torch = self.torch
        stream = torch.cuda.current_stream(self.device)
        call = lambda reps: self._check(self.lib.sdsm_sweep_eval(self.ctx, ctypes.byref(h['plan']), h['d_cands'].data_ptr(), h['n'], int(want_h),
                                                                   int(rows), reps, h['ws'].data_ptr(), stream.cuda_stream))
        call(1)
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        e0.record(stream); call(1); e1.record(stream); call(1 + repeats); e2.record(stream)
        e2.synchronize()
        return (e1.elapsed_time(e2) - e0.elapsed_time(e1)) / repeats       # the prep launch cancels out

    def dfma_peak_gflops(self, iters=20000, repeats=5):
        """Measured register-resident FP64 FMA throughput of this device (GFLOP/s): the FP64 roofline denominator."""
        torch = self.torch
        scratch = torch.zeros(8, dtype=torch.float64, device=self.device)
        flops = ctypes.c_int64()
        best = 0.0
        stream = torch.cuda.current_stream(self.device)
        for _ in range(repeats + 1):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            self._check(self.lib.sdsm_dfma_peak(self.ctx, scratch.data_ptr(), iters, ctypes.byref(flops), stream.cuda_stream))
            e1.record(stream)
            e1.synchronize()
            best = max(best, flops.value / (e0.elapsed_time(e1) * 1e-3) / 1e9)
        return best

    def export_smooth(self, k, n, m):
        """G~ of candidate ``k`` of the last batch: (pixel rc [n], cols [n, 25], vals [n, 25], grid rc [m])."""
        torch = self.torch
        plan, d_cands, ws, cands = self._last
        d_rc = torch.zeros(n, dtype=torch.int32, device=self.device)
        d_cols = torch.full((n * 25,), -1, dtype=torch.int32, device=self.device)
        d_vals = torch.zeros(n * 25, dtype=torch.float32, device=self.device)
        d_grid = torch.zeros(max(m, 1), dtype=torch.int32, device=self.device)
        torch.cuda.synchronize(self.device)
        self._check(self.lib.sdsm_export_smooth(self.ctx, ctypes.byref(plan), d_cands.data_ptr(), k, d_rc.data_ptr(), d_cols.data_ptr(),
                                                d_vals.data_ptr(), d_grid.data_ptr(), ws.data_ptr(), self.stream.cuda_stream))
        self.stream.synchronize()
        return (d_rc.cpu().numpy().view(np.uint32), d_cols.cpu().numpy().reshape(n, 25), d_vals.cpu().numpy().reshape(n, 25),
                d_grid.cpu().numpy().view(np.uint32)[:m])
