"""Batch mode: several images through the hot path together (reference ``superdsm/batch.py:288-361``: one task = one image, tasks
run one after the other or in separate processes).

The stages that call the hot path are sequential per image: the GEM stage calls ``compute_objects`` once for the atoms, once
for the universes and once per generation, and needs the energies of one call before it can enumerate the next
(``globalenergymin.py:183-271,357-364``).  Late generations are small, so one image alone cannot fill a B200.
:class:`LockstepBatch` runs K image jobs -- ordinary Python callables that use the seam, e.g. the reference's unmodified
``Pipeline.process_image`` with ``compute_objects`` rebound -- in K threads and merges their pending
``compute_objects`` calls: whenever every live job is blocked in the seam, the calls become ONE launch over the images of
all jobs (``sdsm_set_images`` / ``sdsm_cand.img``), and the jobs continue with their results.  Nothing in the jobs changes;
the results are bit-identical to running each job alone (every candidate's fit is independent of its batch).

Result formats: :func:`object_rows` / :func:`write_label_csv` produce the (size, centroid) table the reference's regression
suite compares label maps with (``tests/regression/validate.py:26-45``), :func:`rasterize_labels` a label map.
"""
import csv
import threading

import numpy as np


class _Call(object):
    __slots__ = ('job', 'objects', 'y', 'atoms', 'cfg', 'done', 'fallbacks', 'error')


class LockstepBatch(object):
    """``run(jobs)``: ``jobs`` are callables ``job(compute_objects)``; each calls the ``compute_objects`` it is handed exactly as it
    would call ``superdsm.objects.compute_objects`` (same arguments), any number of times, and returns its result."""

    def __init__(self, engine=None, solve=None, max_images=32):
        self.engine = engine
        self._solve = solve                 # solve(calls) -> None, sets the objects; default: the CUDA path
        self.max_images = max_images
        self.launches = 0
        self.calls = 0
        self._images = {}                   # job index -> (content key, DeviceImage)

    # -- the merged launch ------------------------------------------------------------------------------------------------
    def _solve_cuda(self, calls):
        from . import objects as _objects
        eng = self.engine or _objects.get_engine()
        cfg = dict(calls[0].cfg)
        cfg.pop('smooth_mat_max_allocations', None)
        _objects._check_cfg(cfg)
        handles = []
        for c in calls:
            key = _objects._digest(c.y.model, c.y.mask, c.atoms)
            cached = self._images.get(c.job)
            if cached is None or cached[0] != key:
                image = eng.upload_image(c.y.model, c.y.mask, c.atoms, cfg['background_margin'], image=cached[1] if cached else None)
                self._images[c.job] = cached = (key, image)
            handles.append(cached[1])
        eng.bind_images(handles)
        _objects.forget_images()
        eng.set_params(cfg, skip_single_fg=True)
        res = eng.solve_multi([[o.footprint for o in c.objects] for c in calls], borrow=True)
        states, k = [], 0
        for c in calls:
            for o in c.objects:
                states.append((c, o, _objects._state(res, k, o.footprint, res.elapsed / max(1, len(res)))))
                k += 1
        for c, o, (state, fb) in states:
            o.set(state)
            c.fallbacks += int(fb)

    # -- the rendezvous ---------------------------------------------------------------------------------------------------
    def run(self, jobs):
        jobs = list(jobs)
        cond = threading.Condition()
        pending, results, errors = [], [None] * len(jobs), [None] * len(jobs)
        live = [len(jobs)]

        def make_seam(j):
            def compute_objects(objects, y, atoms, dsm_cfg, log_root_dir=None, status_line=None, out=None):
                call = _Call()
                call.job, call.objects, call.y, call.atoms, call.cfg = j, list(objects), y, atoms, dict(dsm_cfg)
                call.done, call.fallbacks, call.error = threading.Event(), 0, None
                if len(call.objects) == 0: return 0
                with cond:
                    pending.append(call)
                    cond.notify_all()
                call.done.wait()
                if call.error is not None: raise call.error
                if out is not None and hasattr(out, 'write') and status_line is not None:
                    out.write(f'{status_line[1]}: {len(call.objects)} ({call.fallbacks}x fallback)')
                return call.fallbacks
            return compute_objects

        def worker(j):
            try: results[j] = jobs[j](make_seam(j))
            except BaseException as err: errors[j] = err
            finally:
                with cond:
                    live[0] -= 1
                    cond.notify_all()

        threads = [threading.Thread(target=worker, args=(j,), daemon=True) for j in range(len(jobs))]
        for t in threads: t.start()
        solve = self._solve or self._solve_cuda
        while True:
            with cond:
                cond.wait_for(lambda: live[0] == 0 or (len(pending) > 0 and len(pending) == live[0]))
                if live[0] == 0 and not pending: break
                calls = list(pending)
                del pending[:]
            # calls with the same hyper-parameters go together, at most max_images per launch
            groups = {}
            for c in calls:
                key = tuple(sorted((k, str(v)) for k, v in c.cfg.items()))
                groups.setdefault(key, []).append(c)
            for group in groups.values():
                for lo in range(0, len(group), self.max_images):
                    part = group[lo:lo + self.max_images]
                    try:
                        solve(part)
                        self.launches += 1
                        self.calls += len(part)
                    except BaseException as err:
                        for c in part: c.error = err
            for c in calls: c.done.set()
        for t in threads: t.join()
        for err in errors:
            if err is not None: raise err
        return results


# ---- result formats ---------------------------------------------------------------------------------------------------------

def object_rows(objects):
    """(size, centre x, centre y) of every object's foreground, sorted: the rows ``tests/regression/validate.py:33-39`` derives from
    a label map (centres rounded to one decimal)."""
    rows = []
    for o in objects:
        if o.fg_fragment is None: continue
        frag = np.asarray(o.fg_fragment, bool)
        if not frag.any(): continue
        rr, cc = np.nonzero(frag)
        rows.append((int(frag.sum()), round(float(cc.mean() + o.fg_offset[1]), 1), round(float(rr.mean() + o.fg_offset[0]), 1)))
    return sorted(rows, key=lambda r: (str(r[1]), str(r[2])))


def write_label_csv(path, rows):
    """The CSV of ``tests/regression/validate.py:40-43``: header + one quoted row per object."""
    with open(path, 'w', newline='') as fp:
        writer = csv.writer(fp, delimiter=',', quoting=csv.QUOTE_ALL)
        writer.writerows([['Object size', 'Center X', 'Center Y']] + [[str(r[0]), str(r[1]), str(r[2])] for r in rows])


def rasterize_labels(shape, objects, background_label=0):
    """uint16 label map of the objects, labels 1.. in order.  The sets of a cover may share atoms, so foregrounds can overlap:
    a pixel claimed by several objects gets the label of the nearest pixel claimed by exactly one (the reference resolves such
    pixels by a watershed of the same distance map, ``render.py:427-433``)."""
    import scipy.ndimage as ndi
    out = np.zeros(shape, np.uint16)
    claims = np.zeros(shape, np.uint8)
    label = 0
    for o in objects:
        if o.fg_fragment is None or not np.asarray(o.fg_fragment).any(): continue
        label += 1
        r, c = (int(v) for v in o.fg_offset)
        h, w = o.fg_fragment.shape
        frag = np.asarray(o.fg_fragment, bool)
        out[r:r + h, c:c + w][frag] = label
        claims[r:r + h, c:c + w][frag] += 1
    contested = claims > 1
    if contested.any():
        sure = (claims == 1)
        if sure.any():
            idx = ndi.distance_transform_edt(~sure, return_distances=False, return_indices=True)
            out[contested] = out[idx[0][contested], idx[1][contested]]
    if background_label != 0: out[claims == 0] = background_label
    return out
