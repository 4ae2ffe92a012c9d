"""superdsm_b200: the per-candidate deformable-shape-model fit of BMCV/SuperDSM, B200-native.

The package mirrors the reference's interface for this path only (``objects.compute_objects``, ``objects.cvxprog``,
``objects.Object``, ``image.Image``, ``dsm.DeformableShapeModel``, the ``dsm/*`` configuration keys); the arithmetic
runs in hand-written CUDA kernels for sm_100a behind the C ABI declared in ``include/superdsm_b200.h``.
"""
VERSION = '0.1.0'
