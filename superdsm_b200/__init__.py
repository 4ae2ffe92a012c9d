"""superdsm_b200: the per-candidate deformable-shape-model fit of BMCV/SuperDSM, B200-native.

The package replaces the reference's hot path behind its own seams (``objects.compute_objects`` / ``_compute_objects``,
``objects.cvxprog``, the normalised-energy computer of the c2f stage) and works on whatever ``Image`` / ``Object`` classes the
caller hands in (duck-typed: ``.model`` / ``.mask``; ``.footprint`` / ``.set``); the arithmetic runs in hand-written CUDA
kernels for sm_100a behind the C ABI declared in ``include/superdsm_b200.h``.
"""
VERSION = '0.2.0'
