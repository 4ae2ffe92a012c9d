"""``sdsm_prepare_image_noalloc``: the per-image preparation with the library's built-in scratch (no allocation in the call).
Collected last on purpose.  Compared, through the C-ABI, with ``sdsm_prepare_image`` (caller scratch) on the same inputs --
integer outputs, so the tolerance is equality -- on two calls in a row (the second with another image: the shared scratch
keeps the first call's contents)."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

NOALLOC_MAX_PIXELS = 4 * 1024 * 1024


def _image(seed, H, W):
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[:H, :W]
    y = -0.3 + 0.05 * rng.standard_normal((H, W))
    for _ in range(12):
        r, c, s = rng.uniform(0, H), rng.uniform(0, W), rng.uniform(6, 18)
        y += 1.2 * np.exp(-((yy - r) ** 2 + (xx - c) ** 2) / (2 * s * s))
    atoms = (1 + (yy // 32) * ((W + 31) // 32) + xx // 32).astype(np.int32)
    mask = rng.uniform(size=(H, W)) > 0.02
    return np.ascontiguousarray(y), mask, atoms


def _run(lib, ctx, torch, dev, y, mask, atoms, margin, noalloc):
    H, W = y.shape
    n_labels = int(atoms.max())
    d_y = torch.from_numpy(y).to(dev)
    d_mask = torch.from_numpy(mask.view(np.uint8)).to(dev)
    d_atoms = torch.from_numpy(atoms).to(dev)
    d_out = torch.full((H, W), -7, dtype=torch.int32, device=dev)
    d_count = torch.full((n_labels + 1,), -7, dtype=torch.int64, device=dev)
    d_box = torch.full((n_labels + 1, 4), -7, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    if noalloc:
        torch.cuda.synchronize(dev)
        before = torch.cuda.memory_allocated(dev)
        rc = lib.sdsm_prepare_image_noalloc(ctx, d_y.data_ptr(), d_mask.data_ptr(), d_atoms.data_ptr(), H, W, margin, n_labels,
                                            d_out.data_ptr(), d_count.data_ptr(), d_box.data_ptr(), stream)
        assert torch.cuda.memory_allocated(dev) == before
    else:
        d_scratch = torch.empty(2 * H * W, dtype=torch.uint8, device=dev)
        rc = lib.sdsm_prepare_image(ctx, d_y.data_ptr(), d_mask.data_ptr(), d_atoms.data_ptr(), H, W, margin, n_labels,
                                    d_out.data_ptr(), d_count.data_ptr(), d_box.data_ptr(), d_scratch.data_ptr(), stream)
    assert rc == 0, lib.sdsm_last_error(ctx)
    torch.cuda.synchronize(dev)
    return d_out.cpu().numpy(), d_count.cpu().numpy(), d_box.cpu().numpy()


def test_prepare_image_noalloc_matches_prepare_image():
    import torch
    from superdsm_b200 import _lib
    from superdsm_b200.engine import prepare_image
    lib = _lib.load()
    dev = torch.device('cuda', 0)
    torch.cuda.set_device(dev)
    ctx = ctypes.c_void_p()
    assert lib.sdsm_create(ctypes.byref(ctx), 0) == 0
    try:
        # two calls in a row on the shared scratch: different images, sizes and margins (the second smaller than the first,
        # so stale distances of the first call lie behind and inside its window)
        for seed, (H, W), margin in ((11, (300, 417), 20.0), (12, (190, 233), 7.5), (13, (64, 96), 1e308 * 10)):
            y, mask, atoms = _image(seed, H, W)
            ref = _run(lib, ctx, torch, dev, y, mask, atoms, margin, noalloc=False)
            got = _run(lib, ctx, torch, dev, y, mask, atoms, margin, noalloc=True)
            for a, b in zip(got, ref): assert np.array_equal(a, b)
            # and both against the numpy statement of the same thing
            atoms_v, count, box = prepare_image(y, mask, atoms, np.inf if margin > 1e300 else margin)
            assert np.array_equal(got[0], atoms_v)
            assert np.array_equal(got[1][1:], count[1:])
        # beyond the built-in scratch: refused with a clear error, nothing launched
        H, W = 2048, 2049
        assert H * W > NOALLOC_MAX_PIXELS
        rc = lib.sdsm_prepare_image_noalloc(ctx, 1, 1, 1, H, W, 5.0, 1, 1, 1, 1, None)
        assert rc == 3                                   # SDSM_E_UNSUPPORTED
        assert b'SDSM_NOALLOC_MAX_PIXELS' in lib.sdsm_last_error(ctx)
    finally:
        lib.sdsm_destroy(ctx)
