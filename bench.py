#!/usr/bin/env python
"""Benchmark of the hot path: DSM candidate solves per second (BASELINE.json metric) on the GOWT1-shaped workload.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path, one JSON line on rank 0
    python bench.py --impl reference --gpus N --steps K ...   # the CPU implementation of the path on the host cores

A step = one pass of the hot path (region packing, smooth-matrix construction, elliptical + DSM solves, result
rasterisation) over every candidate of one synthetic 1024 x 1024 image: atoms, cluster universes and the sets of
adjacent atoms the GEM stage enumerates.  The steps cycle through a pool of 4 images (seeds 1..4); with N > 1 rank r starts
at image r, so that at any time the ranks work on different images while every rank does the same work over the timed
region (image-sharded, no collective on the data path; weak scaling); `value` = candidates of all ranks / max-over-ranks
device time.  `--workload batch`: K images per launch; `--workload tiled4k`: one 4096 x 4096 field, candidate-sharded (strong).
"""
import os

# One BLAS / OpenMP thread per process, set BEFORE numpy loads its BLAS: the CPU arm runs one worker process per host core
# (the reference's process-per-candidate model, objects.py:270-281, MKL_NUM_THREADS in examples/*/task.json); a BLAS pool
# sized to all cores inside every worker would oversubscribe the box and halve the CPU arm's rate.
for _var in ('OMP_NUM_THREADS', 'OPENBLAS_NUM_THREADS', 'MKL_NUM_THREADS', 'NUMEXPR_NUM_THREADS'):
    os.environ[_var] = '1'

import argparse                                     # noqa: E402
import json                                         # noqa: E402
import subprocess                                   # noqa: E402
import sys                                          # noqa: E402
import tempfile                                     # noqa: E402
import time                                         # noqa: E402

import numpy as np                                  # noqa: E402

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path: sys.path.insert(0, ROOT)

WORKLOAD = 'gowt1'
WORKLOAD_DESC = 'GOWT1-shaped 1024x1024 synthetic nuclei image (BASELINE config 1): atoms + universes + adjacent-atom sets, AF_scale 42.43'
# the other BASELINE configurations (`--workload` times them; the default stays the config the metric is quoted on)
OTHER_WORKLOADS = {
    'bbbc039': ('BBBC039-shaped 520x696 image, default configuration (BASELINE config 0)', dict()),
    'nih3t3': ('NIH3T3-shaped 1344x1024 image, clustered nuclei, strong inhomogeneity (BASELINE config 2)', dict()),
    'tiled': ('1024x1024 crop of the 4096x4096 tiled field, 1250 nuclei, AF_scale 10 (BASELINE config 4)', dict(shape=(1024, 1024), n=1250)),
    'batch': ('batch of GOWT1-shaped 1024x1024 images (BASELINE config 3, seeds 1000...): image-sharded over the ranks, K images per launch', dict()),
    'tiled4k': ('4096x4096 tiled field, ~20k nuclei, AF_scale 10 (BASELINE config 4): ONE image, candidates sharded over the ranks (LPT), results gathered on the host', dict()),
}


def make_workload(rank, args, seed=None):
    from workloads import synth
    global WORKLOAD_DESC
    if args.workload == 'tiled4k':
        WORKLOAD_DESC = OTHER_WORKLOADS['tiled4k'][0]
        spec = dict(synth.WORKLOADS['tiled4k'])
        if args.field: spec['shape'], spec['n'] = (args.field, args.field), int(20000 * (args.field / 4096.) ** 2)
        g_raw, seeds, _ = synth.make_image(**spec)
        cfg = synth.dsm_config(10, alpha=3.3e-4 * 100)
        y = synth.offset_image(g_raw, sigma2=10)
        y_mask, atoms, atom_cluster, adjacency, clusters, atom_seeds = synth.make_atoms(y, seeds, 2, np.random.default_rng(spec['seed'] + 1), fast=True)
        fps = synth.enumerate_candidates(atom_cluster, adjacency, 3, args.max_candidates)
        return dict(name='tiled4k', y=y, y_mask=y_mask, atoms=atoms, footprints=fps, dsm_cfg=cfg)
    if args.workload == 'batch':
        WORKLOAD_DESC = OTHER_WORKLOADS['batch'][0]
        return synth.make_workload(WORKLOAD, oversegment=args.oversegment, max_set_size=args.max_set_size, seed=seed, max_candidates=args.max_candidates)
    if args.workload != WORKLOAD:
        WORKLOAD_DESC, kw = OTHER_WORKLOADS[args.workload]
        name = 'tiled4k' if args.workload == 'tiled' else args.workload
        base = synth.WORKLOADS[name]['seed']
        return synth.make_workload(name, oversegment=args.oversegment, max_set_size=args.max_set_size, seed=base + rank,
                                   max_candidates=args.max_candidates, **kw)
    return synth.make_workload(WORKLOAD, oversegment=args.oversegment, max_set_size=args.max_set_size, seed=1 if seed is None else seed,
                               max_candidates=args.max_candidates)


def algorithmic_work(res, counts=None):
    """Algorithmic bytes and flops of one solver launch, from SURVEY.md section 8(d) and the evaluation counters the
    kernel reports: per value+gradient evaluation 12 n + 16 (6 + m) + 8 bytes, per Hessian evaluation
    12 n + 8 (6 + m) + 4 (6 + m)(7 + m) bytes, plus the result pass (8 n); flops likewise with k = nnz(G~) / n <= 25."""
    n = res.cands['n'].astype(np.float64)
    m = res.info[:, 0].astype(np.float64)
    nf = res.info[:, 6].astype(np.float64)
    # info[6] counts the evaluations the reference makes.  Not charged: trials the solver decides without evaluating F (their
    # outcome does not depend on F) and accepted trials evaluated in one pass with their Hessian (sdsm_eval_counts).
    if counts is not None: nf = nf - counts[:, 0] - counts[:, 2]
    nh = res.info[:, 7].astype(np.float64)
    N = 6 + m
    kbar = np.where(m > 0, 12.2, 0.0)            # k = 11-14
    bytes_f = 12 * n + 16 * N + 8
    bytes_h = 12 * n + 8 * N + 4 * N * (N + 1)
    flops_f = n * (12 + 2 * kbar + 3) + 2 * n * (6 + kbar)
    flops_h = n * (6 + kbar) * (7 + kbar)
    # a Hessian evaluation that follows an accepted line-search trial reuses the trial's value and gradient (same point): only
    # the first evaluation of each of the <= 3 solves of a candidate computes both (counted as 2 per candidate on average)
    total_bytes = float((nf * bytes_f + nh * bytes_h + 2 * bytes_f + 8 * n).sum())
    total_flops = float((nf * flops_f + nh * flops_h + 2 * flops_f + nh * N ** 3 / 3).sum())
    return total_bytes, total_flops


class ClockSampler(object):
    def __init__(self, index):
        self.path = tempfile.mktemp(suffix='.csv')
        q = 'clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,' \
            'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap'
        try:
            self.proc = subprocess.Popen(['nvidia-smi', f'--id={index}', f'--query-gpu={q}', '--format=csv,noheader,nounits', '-lms', '100'],
                                         stdout=open(self.path, 'w'), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None: return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        self.proc.terminate()
        try: self.proc.wait(timeout=5)
        except Exception: self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for line in open(self.path):
            parts = [p.strip() for p in line.split(',')]
            if len(parts) < 6: continue
            try: sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError: continue
            for name, flag in zip(names, parts[2:6]):
                if flag.lower().startswith('active'): reasons.add(name)
        try: os.unlink(self.path)
        except OSError: pass
        return {'sm_mhz': float(np.median(sm)) if sm else None, 'sm_max_mhz': float(np.max(mx)) if mx else None,
                'reasons': sorted(reasons), 'samples': len(sm)}


# ---------------------------------------------------------------------------------------------------------------
# CPU implementation of the path (the oracle port; the reference itself is pure Python + cvxopt, which is absent)

def _cpu_worker(job):
    y, y_mask, atoms, fp, cfg = job
    from oracle import dsm_oracle as orc
    t0 = time.perf_counter()
    try:
        res = orc.compute_object(y, y_mask, atoms, fp, cfg)
        energy = res['energy']
    except orc.CvxprogError:                        # a legitimate outcome of the reference too (objects.py:351-353)
        energy = float('nan')
    return energy, time.perf_counter() - t0


def _blas_threads():
    """Effective BLAS / OpenMP thread count of a worker process (reported next to the CPU number)."""
    try:
        from threadpoolctl import threadpool_info
        return max([int(p.get('num_threads', 1)) for p in threadpool_info()] or [1])
    except Exception:
        return None


class CpuArm(object):
    """The CPU implementation of the path (the oracle port) on all host cores: one worker process per core, one BLAS thread
    each.  A step times a stratified sample: the candidates are sorted by pixel count and cut into `sample` strata of equal
    size, step k takes member k (mod size) of every stratum, and every measured time stands for its whole stratum, so that
    the figure is (time for ALL candidates of the workload's images on `cores` cores, perfectly packed) -- candidates / that time."""

    def __init__(self, ws, cores, sample):
        import multiprocessing as mp
        self.ws, self.cores = ws, cores
        self.items, npix = [], []
        for wi in ws:
            count = np.bincount(np.asarray(wi['atoms']).reshape(-1))
            for fp in wi['footprints']:
                self.items.append((wi, fp)); npix.append(sum(int(count[a]) for a in fp))
        order = np.argsort(np.array(npix), kind='stable')
        self.strata = [s for s in np.array_split(order, min(sample, len(self.items))) if len(s)]
        self.pool = mp.get_context('fork').Pool(cores)
        self.blas_threads = self.pool.apply(_blas_threads)
        self.n = len(self.items)

    def step(self, k):
        """Returns (estimated seconds for the whole candidate list on all cores, wall seconds of this step, core-seconds spent)."""
        picks = [int(s[k % len(s)]) for s in self.strata][::-1]              # largest first: better packing of the pool
        weights = [len(s) for s in self.strata][::-1]
        jobs = [(self.items[i][0]['y'], self.items[i][0]['y_mask'], self.items[i][0]['atoms'], self.items[i][1], self.items[i][0]['dsm_cfg']) for i in picks]
        t0 = time.perf_counter()
        out = self.pool.map(_cpu_worker, jobs, chunksize=1)
        wall = time.perf_counter() - t0
        core_s = float(np.sum([o[1] for o in out]))
        est_all = float(np.sum([wt * o[1] for wt, o in zip(weights, out)])) / self.cores
        return est_all, wall, core_s

    def describe(self):
        return (f'{len(self.strata)} of {self.n} candidates per step, stratified by pixel count (one per stratum, rotating), '
                f'each weighted by its stratum size: value = candidates / (estimated time for all {self.n} on {self.cores} cores)')

    def close(self):
        self.pool.close()
        self.pool.join()


def pool_size(args):
    """Images of the default workload's pool: step s of rank r takes image (r + s) mod pool, so that over the timed steps every
    rank does the same work on different images at any one time."""
    return max(1, args.pool) if args.workload == WORKLOAD else 1


def make_pool(rank, args):
    if args.workload == WORKLOAD:
        from workloads import synth
        return [synth.make_workload(WORKLOAD, oversegment=args.oversegment, max_set_size=args.max_set_size, seed=1 + i,
                                    max_candidates=args.max_candidates) for i in range(pool_size(args))]
    return [make_workload(rank, args, seed=1000)]


def workload_config(ws, world, args):
    """The keys that define the workload: identical in both arms."""
    mode = {'batch': 'batch', 'tiled4k': 'sharded'}.get(args.workload, 'single')
    P = pool_size(args)
    sharding = {'single': f'image-sharded: one image per rank and step out of a pool of {P} (seeds 1..{P}; step s of rank r takes image (r + s) mod {P}), one launch per image, no collective',
                'batch': f'image-sharded: {args.images} images per rank and launch (seeds 1000 + rank * {args.images} ...), no collective',
                'sharded': 'candidate-sharded: one image replicated on every rank, candidates split by cost (LPT), packed results gathered on the host (gloo), no collective on the data path'}[mode]
    return {'workload': WORKLOAD_DESC, 'candidates_per_image': [len(wi['footprints']) for wi in ws] if len(ws) > 1 else len(ws[0]['footprints']),
            'images_per_step': {'single': world, 'batch': world * args.images, 'sharded': 1}[mode],
            'candidate_list': f'atoms + universes + adjacent-atom sets up to {3 if mode == "sharded" else args.max_set_size} (oversegment {2 if mode == "sharded" else args.oversegment})',
            'sharding': sharding, 'l2': 'GPU arm: flushed between timed iterations (256 MiB write)'}


def run_reference(args):
    rank = int(os.environ.get('RANK', 0))
    if rank != 0: return
    cores = os.cpu_count() or 1
    ws = make_pool(0, args)
    arm = CpuArm(ws, cores, sample=max(16, min(2 * cores, 64)))
    est, walls = [], []
    t_begin = time.perf_counter()
    for step in range(args.warmup + args.steps):
        if step < args.warmup:
            if step == 0: arm.step(10 ** 6)                  # one untimed pass warms the workers (imports, page cache)
            continue
        e, wall, _ = arm.step(step - args.warmup)
        est.append(e); walls.append(wall)
        if time.perf_counter() - t_begin > 200 and len(est) >= 3: break      # time box: a few minutes in all
    arm.close()
    n = arm.n
    value = n * len(est) / float(np.sum(est))
    desc = arm.describe()
    line = {'impl': 'reference', 'metric': 'dsm_candidate_solves_per_s', 'value': value, 'unit': 'candidate-solves/s',
            'n_gpus': args.gpus, 'steps': len(est), 'warmup': args.warmup, 'ms_per_step': 1e3 * float(np.mean(walls)),
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
            'config': workload_config(ws, max(1, args.gpus), args),
            'cpu_baseline': {'value': value, 'unit': 'candidate-solves/s', 'cores': cores, 'blas_threads_per_process': arm.blas_threads,
                             'kind': 'port', 'sample': desc,
                             'note': 'numpy/scipy port of dsm.Energy + restated cvxopt cp (oracle/); the reference needs cvxopt + MKL, unavailable on this box'},
            'e2e': {'value': value, 'unit': 'candidate-solves/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'gpu_launches': 0}
    print(json.dumps(line))


# ---------------------------------------------------------------------------------------------------------------

def run_gpu(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get('RANK', 0))
    local_rank = int(os.environ.get('LOCAL_RANK', 0))
    world = int(os.environ.get('WORLD_SIZE', 1))
    torch.cuda.set_device(local_rank)
    device = torch.device('cuda', local_rank)
    if world > 1:
        dist.init_process_group('nccl', device_id=device)

    from superdsm_b200 import _lib
    from tests.compat.refshim import Image, Object
    from superdsm_b200.objects import compute_objects, compute_objects_multi, forget_images, get_engine
    eng = get_engine(local_rank)
    gloo = dist.new_group(backend='gloo') if world > 1 else None      # host-side result gather of the candidate-sharded mode
    mode = {'batch': 'batch', 'tiled4k': 'sharded'}.get(args.workload, 'single')

    # ---- the units of work of this rank: `units[i]` = (images bound for the launch, prepared handle, footprint lists) ----
    units = []
    if mode == 'batch':
        # K images per rank and launch, distinct seeds over all ranks (BASELINE config 3: seeds 1000 ...)
        ws = [make_workload(rank, args, seed=1000 + rank * args.images + i) for i in range(args.images)]
        images = [eng.upload_image(wi['y'], wi['y_mask'], wi['atoms'], wi['dsm_cfg']['background_margin']) for wi in ws]
        eng.bind_images(images)
        eng.set_params(ws[0]['dsm_cfg'])
        units.append(dict(images=images, ws=ws, handle=eng.prepare([wi['footprints'] for wi in ws], multi=True)))
    elif mode == 'sharded':
        # ONE image on every rank, the candidates split by cost (LPT): strong scaling
        from superdsm_b200.engine import lpt_shards
        from superdsm_b200.sharding import candidate_costs_from_counts
        ws = [make_workload(rank, args)]
        all_fps = ws[0]['footprints']
        image = eng.upload_image(ws[0]['y'], ws[0]['y_mask'], ws[0]['atoms'], ws[0]['dsm_cfg']['background_margin'])
        eng.bind_images([image])
        eng.set_params(ws[0]['dsm_cfg'])
        mine = lpt_shards(candidate_costs_from_counts(all_fps, image.atom_count), world)[rank]
        shard = dict(ws[0]); shard['footprints'] = [all_fps[i] for i in mine]
        units.append(dict(images=[image], ws=[shard], handle=eng.prepare(shard['footprints'])))
    else:
        ws = make_pool(rank, args)
        eng.set_params(ws[0]['dsm_cfg'])
        for wi in ws:
            image = eng.upload_image(wi['y'], wi['y_mask'], wi['atoms'], wi['dsm_cfg']['background_margin'])
            eng.bind_images([image])
            units.append(dict(images=[image], ws=[wi], handle=eng.prepare(wi['footprints'])))
    P = len(units)
    unit_of = lambda step: units[(rank + step) % P]
    stream = torch.cuda.current_stream(device)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)        # > 126 MB L2

    def barrier():
        if world > 1: dist.barrier()
        torch.cuda.synchronize(device)

    def launch(u):
        eng.bind_images(u['images'])                                        # host-side: the image set travels as a kernel argument
        eng.run_prepared(u['handle'], stream)

    # ---- device-resident throughput (`value`) ----
    for s_ in range(args.warmup): launch(unit_of(s_))
    barrier()
    launches0 = eng.launch_count
    sampler = ClockSampler(local_rank)
    events = []
    barrier()
    for s_ in range(args.steps):
        flush.fill_(1)                                                      # L2 flush between timed iterations
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        launch(unit_of(args.warmup + s_))
        e1.record(stream)
        events.append((e0, e1))
    barrier()
    clocks = sampler.stop()
    gpu_launches = eng.launch_count - launches0
    step_ms = [a.elapsed_time(b) for a, b in events]
    total_ms = float(np.sum(step_ms))
    if args.warmup + args.steps < P:
        for u in units: launch(u)                                           # every unit has results to read below
        torch.cuda.synchronize(device)

    # ---- results of every unit (work model, sanity check); CvxprogError candidates (both elliptical passes failed,
    #      objects.py:351-353) are legitimate results of the reference too, but its compute_objects raises on them: they are
    #      dropped from the end-to-end leg ----
    optimal = n_errors = 0
    stats = None
    for u in units:
        h = u['handle']
        stats = eng.solver_stats(h['ws'], h['plan'])
        if stats['abort'] or stats['done'] != h['n']: raise SystemExit(f'persistent solver did not finish: {stats}')
        n = h['n']
        info = h['out']['info'][:32 * n].cpu().numpy().view(np.int32).reshape(n, 8)
        status = h['out']['status'][:4 * n].cpu().numpy().view(np.int32)

        class _R(object): pass
        r = _R(); r.cands = h['cands'].copy(); r.info = info
        counts = eng.eval_counts(h['ws'], n, h['plan'])
        u['alg_bytes'], u['alg_flops'] = algorithmic_work(r, counts)
        u['alg_bytes_ref'] = algorithmic_work(r)[0]          # charged with every evaluation the reference makes
        u['evals'] = np.array([int(info[:, 6].sum()), int(info[:, 7].sum())] + counts.sum(axis=0).tolist(), dtype=np.int64)
        u['n'], u['pixels'] = n, int(h['cands']['n'].sum())
        optimal += int((status == _lib.ST_OPTIMAL).sum()); n_errors += int((status >= _lib.ST_ERROR).sum())
        pos = 0
        u['fps'] = []
        for wi in u['ws']:
            k = len(wi['footprints'])
            u['fps'].append([fp for fp, st in zip(wi['footprints'], status[pos:pos + k]) if st < _lib.ST_ERROR])
            pos += k
    timed_units = [unit_of(args.warmup + s_) for s_ in range(args.steps)]
    cand_local = float(sum(u['n'] for u in timed_units))
    alg_bytes = float(sum(u['alg_bytes'] for u in timed_units)); alg_flops = float(sum(u['alg_flops'] for u in timed_units))
    alg_bytes_ref = float(sum(u['alg_bytes_ref'] for u in timed_units))
    pixels_local = int(np.mean([u['pixels'] for u in timed_units]))

    # ---- end to end through the public API (`e2e`): host arrays in, Objects out ----
    if mode == 'sharded':
        from superdsm_b200.sharding import compute_objects_sharded
        bad = set(fp for fp in units[0]['ws'][0]['footprints'] if fp not in set(units[0]['fps'][0]))
        if world > 1:
            parts = [None] * world
            dist.all_gather_object(parts, bad, group=gloo)
            bad = set().union(*parts)
        keep = [fp for fp in all_fps if fp not in bad]
        w0 = ws[0]
        for u in units: del u['handle']
        torch.cuda.empty_cache()
        y_img = Image(w0['y'], w0['y_mask'])

        def e2e_step(s_):
            objs = [Object(fp) for fp in keep]
            forget_images()
            compute_objects_sharded(objs, y_img, w0['atoms'], w0['dsm_cfg'], None, group=gloo)     # every rank ends with all results
            return objs, len(keep) / world
    elif mode == 'batch':
        def e2e_step(s_):
            u = units[0]
            jobs = [([Object(fp) for fp in fps], Image(wi['y'], wi['y_mask']), wi['atoms']) for wi, fps in zip(u['ws'], u['fps'])]
            compute_objects_multi(jobs, u['ws'][0]['dsm_cfg'])          # uploads every image, one launch, Objects out
            return [ob for objs, _, _ in jobs for ob in objs], sum(len(fps) for fps in u['fps'])
    else:
        def e2e_step(s_):
            u = unit_of(s_)
            wi = u['ws'][0]
            objs = [Object(fp) for fp in u['fps'][0]]
            forget_images()                                                # every step uploads its image again
            compute_objects(objs, Image(wi['y'], wi['y_mask']), wi['atoms'], wi['dsm_cfg'], None)
            return objs, len(objs)
    e2e_step(0)
    h2d0, d2h0 = eng.h2d_bytes, eng.d2h_bytes
    barrier()
    t0 = time.perf_counter()
    e2e_steps = max(1, min(args.steps, 2 * P if mode == 'single' else 3))
    e2e_cands = 0.0
    for s_ in range(e2e_steps):
        objs, cnt = e2e_step(s_)
        e2e_cands += cnt
    torch.cuda.synchronize(device)
    e2e_s = time.perf_counter() - t0
    h2d = (eng.h2d_bytes - h2d0) // e2e_steps
    d2h = (eng.d2h_bytes - d2h0) // e2e_steps
    assert all(np.isfinite(ob.energy) for ob in objs)

    # ---- max over ranks (times), sum over ranks (units) ----
    n_images = {'single': 1, 'batch': args.images, 'sharded': 1.0 / world}[mode]
    t = torch.tensor([total_ms, e2e_s * 1e3, cand_local, e2e_cands, float(n_images)], dtype=torch.float64, device=device)
    if world > 1:
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        tmin = t.clone(); dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
        total_ms_max, e2e_ms_max, cand_total, e2e_cand_total, img_per_step = float(tmax[0]), float(tmax[1]), float(tsum[2]), float(tsum[3]), float(tsum[4])
        total_ms_min = float(tmin[0])
    else:
        total_ms_max, e2e_ms_max, cand_total, e2e_cand_total, img_per_step = total_ms, e2e_s * 1e3, cand_local, e2e_cands, float(n_images)
        total_ms_min = total_ms

    if rank == 0:
        peaks = {}
        try: peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
        except Exception: pass
        hbm_peak = float(peaks.get('hbm_gbs', 6650.0))
        peak_src = 'MEASURED_PEAKS.json hbm_gbs (of measured)' if 'hbm_gbs' in peaks else '6650 GB/s (of fallback)'
        dfma = eng.dfma_peak_gflops()
        traffic, traffic_src = None, None          # dram bytes of one k_solve launch, from the last `ncu --set full` capture
        try:
            tj = json.load(open(os.path.join(ROOT, 'profiles', 'ksolve_traffic.json')))
            traffic, traffic_src = float(tj['dram_bytes_per_launch']), tj['source']
        except Exception: pass
        ms_per_step = total_ms_max / args.steps
        achieved = alg_bytes / (total_ms * 1e-3) / 1e9   # rank 0's launches: the solver kernel is > 97 % of a step (profiles/), both on this stream
        value = cand_total / (total_ms_max * 1e-3)
        line = {
            'metric': 'dsm_candidate_solves_per_s', 'value': value, 'unit': 'candidate-solves/s', 'n_gpus': world, 'steps': args.steps,
            'warmup': args.warmup, 'ms_per_step': ms_per_step, 'higher_is_better': True, 'scaling': 'strong' if mode == 'sharded' else 'weak', 'vs_baseline': None,
            'dtype': 'f64', 'data': 'synthetic',
            'config': workload_config(ws, world, args),
            'stats': {'pixels_per_step_per_gpu': pixels_local, 'optimal': optimal, 'errors': n_errors, 'candidates_per_step': cand_total / args.steps,
                      'images_per_step': img_per_step, 'images_per_s': img_per_step * args.steps / (total_ms_max * 1e-3),
                      'e2e_images_per_s': img_per_step * e2e_steps / (e2e_ms_max * 1e-3),
                      'rank0_step_ms': [round(v, 2) for v in step_ms], 'slowest_over_fastest_rank': total_ms_max / max(total_ms_min, 1e-9)},
            'e2e': {'value': e2e_cand_total / (e2e_ms_max * 1e-3), 'unit': 'candidate-solves/s', 'h2d_bytes_per_step': int(h2d),
                    'd2h_bytes_per_step': int(d2h), 'includes': 'pinned H2D of y, mask, atoms; device EDT + atom statistics; descriptors; solve; D2H of results; Object materialisation' + ('; result gather over the ranks' if mode == 'sharded' else '')},
            'gpu_launches': int(gpu_launches),
            'roofline': {'kernel': 'k_solve (persistent solver: control CTAs + evaluation warps fed by device-side priority queues)', 'bound': 'hbm', 'achieved': achieved, 'peak': hbm_peak,
                         'unit': 'GB/s', 'frac': achieved / hbm_peak, 'traffic': traffic, 'traffic_source': traffic_src, 'peak_source': peak_src,
                         'algorithmic_bytes_per_launch': alg_bytes / args.steps,
                         'charged': 'evaluations executed and needed (sdsm_eval_counts): not the line-search trials decided without evaluating F, one pass for a trial evaluated with its Hessian',
                         'frac_if_charged_with_every_reference_evaluation': alg_bytes_ref / (total_ms * 1e-3) / 1e9 / hbm_peak,
                         'fp64': {'achieved_gflops': alg_flops / (total_ms * 1e-3) / 1e9, 'peak_gflops': dfma,
                                  'frac': alg_flops / (total_ms * 1e-3) / 1e9 / dfma, 'peak_source': 'register-resident DFMA loop measured in this run',
                                  'note': 'the kernel is bound by FP64 / conversion issue and dependent-issue latency at 16 warps per SM, not by HBM (DESIGN.md section 4)'}},
            'clocks': clocks,
            'solver': {k: stats[k] for k in ('n_eval', 'n_hrows', 'n_ctrl', 'n_promoted')} | {
                'evaluations_all_units': dict(zip(('reference_value_gradient', 'reference_hessian', 'decided_without_evaluation', 'evaluated_ahead_unused',
                                                   'merged_with_hessian', 'hessians_in_vain'), sum(u['evals'] for u in units).tolist())),
                'control_cta_time_share': {k[4:]: round(stats[k] / max(stats['cyc_total'], 1), 4) for k in ('cyc_ctrl', 'cyc_idle')},
                'evaluation_warp_time_share': {k[4:]: round(stats[k] / max(stats['cyc_wtotal'], 1), 4) for k in ('cyc_eval', 'cyc_hrows', 'cyc_widle')}},
        }
        if world == 1 and not args.no_cpu_baseline:
            cores = os.cpu_count() or 1
            arm = CpuArm(ws, cores, sample=max(16, min(2 * cores, 64)))
            est_all, wall, cpu_s = arm.step(0)
            arm.close()
            line['cpu_baseline'] = {'value': arm.n / est_all, 'unit': 'candidate-solves/s', 'cores': cores,
                                    'blas_threads_per_process': arm.blas_threads, 'kind': 'port',
                                    'sample': arm.describe() + f' ({wall:.1f} s wall, {cpu_s:.1f} core-s)',
                                    'note': 'numpy/scipy port of dsm.Energy + restated cvxopt cp (oracle/); cvxopt and MKL are unavailable on this box'}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=10)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--oversegment', type=int, default=3)
    ap.add_argument('--max-set-size', type=int, default=4)
    ap.add_argument('--max-candidates', type=int, default=None)
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--images', type=int, default=8, help='--workload batch: images per rank and launch')
    ap.add_argument('--pool', type=int, default=4, help='default workload: images in the pool the steps cycle through')
    ap.add_argument('--field', type=int, default=0, help='--workload tiled4k: edge of the field in pixels (default 4096)')
    ap.add_argument('--workload', default=WORKLOAD, choices=[WORKLOAD] + sorted(OTHER_WORKLOADS))
    args = ap.parse_args()
    if args.impl == 'reference': run_reference(args)
    else: run_gpu(args)


if __name__ == '__main__':
    main()
