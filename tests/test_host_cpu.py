"""CPU-only checks of the host logic and of the C-ABI library (no compute calls)."""
import ctypes
import os
import re

import numpy as np
import pytest

from superdsm_b200 import _lib, engine, objects, dsm
from tests.compat import refshim
from tests.golden_util import CASES, GoldenCase

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, 'include', 'superdsm_b200.h')).read()
    declared = set(re.findall(r'\b(sdsm_[a-z_]+)\s*\(', header))
    assert declared == set(_lib.EXPORTS)
    lib = _lib.load()
    for name in declared: assert hasattr(lib, name), name
    assert lib.sdsm_abi_version() == _lib.ABI_VERSION == 5


def test_struct_layouts_match_the_header():
    assert _lib.CAND_DTYPE.itemsize == 96 and _lib.CAND_DTYPE.fields['pix_off'][1] == 40 and _lib.CAND_DTYPE.fields['img'][1] == 32
    assert ctypes.sizeof(_lib.Params) == 96 and ctypes.sizeof(_lib.Plan) == 96 and ctypes.sizeof(_lib.Out) == 64
    assert ctypes.sizeof(_lib.ImageDesc) == 24
    # the header's own sizes, from a C compiler
    import subprocess, tempfile
    src = '#include <stdio.h>\n#include "superdsm_b200.h"\nint main(){printf("%zu %zu %zu %zu %zu", sizeof(sdsm_cand), sizeof(sdsm_params), sizeof(sdsm_plan), sizeof(sdsm_out), sizeof(sdsm_image));return 0;}'
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, 't.c'), 'w').write(src)
        subprocess.run(['gcc', '-I', os.path.join(ROOT, 'include'), '-o', os.path.join(d, 't'), os.path.join(d, 't.c')], check=True)
        sizes = subprocess.run([os.path.join(d, 't')], capture_output=True, text=True).stdout.split()
    assert [int(v) for v in sizes] == [96, 96, 96, 64, 24]


@pytest.mark.parametrize('name', CASES)
def test_host_region_descriptors(name):
    """prepare_image + describe_candidates reproduce Object.get_cvxprog_region (objects.py:95-128) exactly."""
    c = GoldenCase(name)
    atoms_v, count, box = engine.prepare_image(c.y, c.y_mask, c.atoms, c.dsm_cfg['background_margin'])
    cands, flat = engine.describe_candidates(count, box, c.footprints)
    for k in range(len(c)):
        mask = c.mask(k)
        assert np.array_equal(np.isin(atoms_v, sorted(c.footprints[k])), mask)
        rows, cols = np.nonzero(mask.any(axis=1))[0], np.nonzero(mask.any(axis=0))[0]
        assert int(cands['n'][k]) == mask.sum()
        assert (cands['r0'][k], cands['h'][k]) == (rows[0], rows[-1] - rows[0] + 1)
        assert (cands['c0'][k], cands['w'][k]) == (cols[0], cols[-1] - cols[0] + 1)


def test_describe_rejects_bad_footprints():
    c = GoldenCase('hand')
    _, count, box = engine.prepare_image(c.y, c.y_mask, c.atoms, 9)
    with pytest.raises(engine.SolverError): engine.describe_candidates(count, box, [{99}])
    with pytest.raises(engine.SolverError): engine.describe_candidates(count, box, [set()])


def test_psf_quarter_table_is_the_reference_kernel():
    import scipy.ndimage as ndi
    q, R = engine.gaussian_psf32(4, 2)
    assert R == 16 and q.shape == (17, 17) and q.dtype == np.float32
    inp = np.zeros((33, 33)); inp[16, 16] = 1
    full = ndi.gaussian_filter(inp, 4).astype(np.float32)
    for dr in (-16, -3, 0, 5, 16):
        for dc in (-16, 0, 7): assert full[16 + dr, 16 + dc] == q[abs(dr), abs(dc)]
    with pytest.raises(AssertionError): engine.gaussian_psf32(3.3, 1)        # even box: dsm.py:147 asserts


def test_lpt_shards_balance_and_cover():
    rng = np.random.default_rng(0)
    costs = rng.integers(1, 1000, 500).astype(float)
    shards = engine.lpt_shards(costs, 8)
    assert sorted(np.concatenate(shards).tolist()) == list(range(500))
    loads = np.array([costs[s].sum() for s in shards])
    assert loads.max() / loads.mean() < 1.02


def test_caller_side_shim_matches_the_reference_fixtures():
    """tests/test_objects.py and tests/test_image.py of the reference on the caller-side stand-ins the GPU tests use
    (tests/compat/refshim.py), and the light model container the seam returns."""
    obj = refshim.Object()
    obj.fg_fragment = np.array([[False, True], [True, True], [True, False]])
    obj.fg_offset = (1, 2)
    actual = np.zeros((4, 5), bool)
    obj.fill_foreground(actual)
    expected = np.array([[0, 0, 0, 0, 0], [0, 0, 0, 1, 0], [0, 0, 1, 1, 0], [0, 0, 1, 0, 0]], bool)
    assert np.array_equal(actual, expected)
    o = refshim.Object({2, 3})
    assert np.array_equal(o.get_mask(np.array([[1, 1, 2], [1, 3, 2], [3, 3, 3]])), np.array([[0, 0, 1], [0, 1, 1], [1, 1, 1]], bool))
    y = np.array([[-1] * 5] * 3 + [[-1, 1, -1, -1, -1], [-1, 1, -1, -1, 1], [-1, 1, -1, -1, 1]])
    atoms = np.array([[1] * 5, [1] * 5, [1, 1, 1, 1, 2], [1, 1, 1, 2, 2], [1, 1, 1, 2, 2], [1, 1, 1, 2, 2]])
    region = refshim.Object({1}).get_cvxprog_region(refshim.Image(y), atoms, background_margin=2)
    assert np.array_equal(region.mask, np.array([[0] * 5, [0, 1, 0, 0, 0]] + [[1, 1, 1, 0, 0]] * 4, bool))     # tests/test_objects.py:50-74
    big = np.zeros(y.shape, bool); big[1:4, 2:4] = expected[1:4, 2:4]
    crop = refshim.Image(y.astype(float)).get_region(big, shrink=True)
    assert crop.offset == (1, 2) and crop.mask.shape == (3, 2) and crop.full_mask.shape == y.shape
    m = dsm.ModelParams(np.arange(8.))
    assert np.array_equal(m.A, [[0, 2], [2, 1]]) and np.array_equal(m.ξ, [6, 7]) and m.c == 5
    with pytest.raises(ValueError): dsm.ModelParams(np.arange(5.))
    st = objects.ObjectState()
    st.footprint, st.energy, st.on_boundary, st.is_optimal, st.processing_time = {4}, 1.5, False, True, 0.1
    st.fg_offset, st.fg_fragment = np.array([1, 2]), np.ones((2, 2), bool)
    o = refshim.Object().set(st)
    assert o.footprint == {4} and o.energy == 1.5 and o.fg_fragment.shape == (2, 2)


def test_engine_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available(): pytest.skip('GPU present')
    with pytest.raises(engine.SolverError, match='no CPU fallback'): engine.Engine(0)


def _gloo_worker(rank, world, port, out_path):
    import os
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from superdsm_b200 import sharding
    rng = np.random.default_rng(7)                       # same data on both ranks
    atoms = rng.integers(0, 40, (64, 64)).astype(np.int32)
    fps = [set(rng.choice(np.arange(1, 40), size=int(rng.integers(1, 4)), replace=False).tolist()) for _ in range(57)]
    objs = []
    for fp in fps:
        o = refshim.Object(fp); objs.append(o)
    solved_here = []

    def fake_solve(shard, y, atoms_, cfg, log):        # stands in for the CUDA engine: a state that depends on the footprint only
        for k, o in enumerate(shard):
            st = refshim.Object(o.footprint)
            st.energy = float(sum(o.footprint)) * cfg['scale']; st.on_boundary = False; st.is_optimal = True; st.processing_time = 0.0
            st.fg_offset = np.array([min(o.footprint), 0]); st.fg_fragment = np.ones((1, len(o.footprint)), bool)
            solved_here.append(sorted(o.footprint))
            yield k, st, (min(o.footprint) % 5 == 0)
    fb = sharding.compute_objects_sharded(objs, None, atoms, dict(scale=2.0), None, solve_shard=fake_solve)
    ok = all(o.energy == 2.0 * sum(o.footprint) and o.fg_fragment.shape[1] == len(o.footprint) for o in objs)
    np.save(out_path % rank, np.array([int(ok), fb, len(solved_here), len(objs)]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_sharding(tmp_path):
    """world_size 2 on gloo: every candidate is solved by exactly one rank and every rank ends up with all results."""
    import socket
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0)); port = s.getsockname()[1]
    out = str(tmp_path / 'r%d.npy')
    mp.spawn(_gloo_worker, args=(2, port, out), nprocs=2, join=True)
    r0, r1 = np.load(out % 0), np.load(out % 1)
    assert r0[0] == 1 and r1[0] == 1                     # all objects computed on both ranks
    assert r0[1] == r1[1] and r0[1] > 0                   # same fallback count everywhere
    assert r0[2] + r1[2] == r0[3] and min(r0[2], r1[2]) > 0    # disjoint shards covering the batch


def test_logistic_restatement_accuracy(tmp_path):
    """csrc/logistic.cuh (the device's exp / log(1 + h) / h / (1 + h), written for this path instead of the CUDA math
    library) compiled for the host and compared with glibc in extended precision: exp <= 1.05 ulp, log <= 1 ulp, the
    division correctly rounded; overflow branch and special values of dsm.py:298-322."""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / 'logistic_check')
    subprocess.run(['g++', '-O2', '-ffp-contract=off', '-o', exe, os.path.join(root, 'tests', 'native', 'logistic_check.cpp')], check=True)
    res = subprocess.run([exe], capture_output=True, text=True)
    assert res.returncode == 0, res.stdout + res.stderr
    assert 'max ulps' in res.stdout
