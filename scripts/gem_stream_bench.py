"""Throughput on the reference's REAL candidate stream (tests/golden/gem_gowt1.npz: the calls the unmodified GEM stage makes on the
GOWT1-shaped image, 2 008 candidates in 11 dependent compute_objects calls): K copies of the job in lock-step
(superdsm_b200.batch.LockstepBatch), end to end through the seam (host arrays in, objects out)."""
import sys, time, json
sys.path.insert(0, '.')
import numpy as np
from superdsm_b200.batch import LockstepBatch
from tests.test_batch_driver import _stream_job

out = []
for K in [int(v) for v in sys.argv[1:]] or [1, 2, 4, 8]:
    sink = {}
    jobs = [_stream_job('gowt1', sink)[0] for _ in range(K)]
    lb = LockstepBatch()
    lb.run(jobs)                                   # warm-up (uploads, allocations)
    t0 = time.perf_counter()
    reps = 2
    for _ in range(reps): counts = lb.run(jobs)
    dt = (time.perf_counter() - t0) / reps
    out.append(dict(jobs=K, candidates=int(sum(counts)), seconds=round(dt, 4), candidate_solves_per_s=round(sum(counts) / dt, 1), images_per_s=round(K / dt, 3),
                    launches_per_run=lb.launches // (reps + 1)))
    print(json.dumps(out[-1]))
