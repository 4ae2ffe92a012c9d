"""The hot path of the reference's coarse-to-fine region analysis, batched (``superdsm/c2freganal.py:58-79,193-287``).

``C2F_RegionAnalysis`` decides every cluster split by comparing normalised energies ``J(model) / n`` of 6-parameter fits
(``smooth_amount = inf``, ``c2freganal.py:126``).  The reference computes them one at a time inside every cluster's own
loop (two fits per split attempt, ``c2freganal.py:205,252``).  Here

* :class:`NormalizedEnergyComputer` fits MANY regions of one image in one launch: the image is uploaded once, the regions of a
  round travel as one int32 label map (regions of different clusters -- and the two halves of a split -- are disjoint), the
  fits run as one batch of the persistent solver with no deformation parameters;
* :func:`run_level_synchronous` advances every cluster's split state machine in lock-step: the state machines are generators
  (the reference's ``_process_cluster_impl`` with ``compute_normalized_energy(...)`` turned into ``yield``, see
  ``INTEGRATION.md``) that yield the regions they need next and are sent the energies.

Seeds, watershed and the split decisions stay host code of the caller, unchanged.
"""
import hashlib

import numpy as np

from . import _lib
from .engine import SolverError
from .objects import CvxprogError, get_engine


class RegionRequest(object):
    """One region whose normalised energy is wanted: ``mask`` (bool) over the window ``offset`` + ``mask.shape`` of the image.
    For the c2f stage: ``cp_region.mask`` of ``Object.get_cvxprog_region`` on the cluster crop and the crop's offset."""
    __slots__ = ('offset', 'mask')

    def __init__(self, offset, mask):
        self.offset = (int(offset[0]), int(offset[1]))
        self.mask = np.asarray(mask, dtype=bool)

    def key(self):
        return (self.offset, self.mask.shape, hashlib.sha1(np.packbits(self.mask)).digest())


class NormalizedEnergyComputer(object):
    """``compute_normalized_energy`` of ``c2freganal.py:58-79`` for batches of regions of ONE image.

    Conventions of the reference: a region whose offset values are all positive or all negative has no energy (``None``);
    results are cached by the region mask; the fit uses ``smooth_amount = inf`` whatever ``dsm_cfg`` says
    (``c2freganal.py:126``); the energy is ``J(model) / n`` in the full-image frame (``c2freganal.py:71-75``).
    """

    def __init__(self, y_model, dsm_cfg, engine=None):
        self.engine = engine or get_engine()
        self.y = np.ascontiguousarray(y_model, dtype=np.float64)
        cfg = dict(dsm_cfg)
        cfg.pop('smooth_mat_max_allocations', None)
        cfg.pop('background_margin', None)           # already applied by the caller (get_cvxprog_region on the cluster crop)
        cfg['smooth_amount'] = np.inf
        self.cfg = cfg
        self.cache = {}
        self.solved = {}                             # key -> (request, cvxopt status bits of the fit): diagnostics
        self.image = None
        self.uploads = 0                             # uploads of y (stays 1: the rounds only send label maps)
        self.launches = 0                            # solver batches

    def _solve_layer(self, reqs):
        """One label map, one batch.  `reqs`: list of (index, RegionRequest), pairwise disjoint."""
        eng = self.engine
        H, W = self.y.shape
        labels = np.zeros((H, W), np.int32)
        for lab, (_, rq) in enumerate(reqs, start=1):
            (r, c), (h, w) = rq.offset, rq.mask.shape
            labels[r:r + h, c:c + w][rq.mask] = lab
        ones = np.ones((H, W), bool)
        if self.image is None:
            self.image = eng.upload_image(self.y, ones, labels, np.inf)
            self.uploads += 1
        else:
            eng.update_labels(self.image, ones, labels, np.inf)
        eng.bind_images([self.image])
        eng.set_params(self.cfg, skip_single_fg=False)
        res = eng.solve([{lab} for lab in range(1, len(reqs) + 1)])
        self.launches += 1
        out = []
        for k in range(len(reqs)):
            st = int(res.status[k])
            if st == _lib.ST_ERROR: out.append(CvxprogError('elliptical initialisation failed'))
            elif st == _lib.ST_BADREGION: raise SolverError('region mismatch between host and device')
            else: out.append(float(res.energy[k]) / float(res.cands['n'][k]))
            self.solved[reqs[k][1].key()] = (reqs[k][1], int(res.info[k, 5]), res.info[k, 1:4].copy())
        return out

    def compute(self, requests):
        """Normalised energies of the requested regions, in order: a float, ``None`` (region all-positive or all-negative,
        ``c2freganal.py:68-69``), or a :class:`CvxprogError` instance where the reference's ``cvxprog`` would have raised."""
        requests = list(requests)
        results = [None] * len(requests)
        todo = []
        seen = {}
        for i, rq in enumerate(requests):
            key = rq.key()
            if key in self.cache: results[i] = self.cache[key]; continue
            if key in seen: continue
            (r, c), (h, w) = rq.offset, rq.mask.shape
            values = self.y[r:r + h, c:c + w][rq.mask]
            if len(values) == 0 or (values > 0).all() or (values < 0).all():
                self.cache[key] = None
                continue
            seen[key] = i
            todo.append((i, rq, key))
        # layers of pairwise disjoint regions (one layer in the c2f stage: clusters are disjoint, and so are the halves of a split)
        while todo:
            taken = np.zeros(self.y.shape, bool)
            layer, rest = [], []
            for item in todo:
                i, rq, key = item
                (r, c), (h, w) = rq.offset, rq.mask.shape
                win = taken[r:r + h, c:c + w]
                if (win & rq.mask).any(): rest.append(item)
                else:
                    win |= rq.mask
                    layer.append(item)
            for (i, rq, key), e in zip(layer, self._solve_layer([(i, rq) for i, rq, _ in layer])):
                self.cache[key] = e
            todo = rest
        for i, rq in enumerate(requests):
            results[i] = self.cache[rq.key()]
        return results


def run_level_synchronous(tasks, computer):
    """Advances the split state machines of all clusters of an image in lock-step.

    ``tasks``: iterable of generators.  A generator yields a list of :class:`RegionRequest` (the regions whose normalised
    energies it needs next: the root of a cluster, or the two halves of a split attempt) and is sent the list of results of
    :meth:`NormalizedEnergyComputer.compute`; its return value is the cluster's result.  Every round makes ONE call of
    ``computer.compute`` with the requests of all live clusters.  Returns the list of return values, in task order.
    """
    gens = list(tasks)
    results = [None] * len(gens)
    pending = {}
    for i, g in enumerate(gens):
        try: pending[i] = list(next(g))
        except StopIteration as stop: results[i] = stop.value
    rounds = 0
    while pending:
        order = sorted(pending)
        flat = [rq for i in order for rq in pending[i]]
        energies = computer.compute(flat)
        rounds += 1
        pos = 0
        nxt = {}
        for i in order:
            cnt = len(pending[i])
            try: nxt[i] = list(gens[i].send(energies[pos:pos + cnt]))
            except StopIteration as stop: results[i] = stop.value
            pos += cnt
        pending = nxt
    run_level_synchronous.last_rounds = rounds
    return results
