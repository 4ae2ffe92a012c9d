"""Host-side mirror of the parts of the reference's ``superdsm/dsm.py`` that callers of the hot path touch.

The arithmetic of ``dsm.Energy`` / ``dsm.CP`` lives in the CUDA library; what remains here is the parameter container
``DeformableShapeModel`` (reference ``dsm.py:22-128``) and a light ``Energy`` handle returned by ``objects.cvxprog``.
"""
import numpy as np


class DeformableShapeModel:
    """Parameters (theta, xi) of a deformable shape model: ``array = [a1, a2, a3, b1, b2, c, xi...]`` (dsm.py:48-63)."""

    def __init__(self, *args):
        if len(args) == 1 and len(args[0]) >= 6:
            self.array = np.asarray(args[0]).astype(float).reshape(-1)
            self.a = self.array[:3]
            self.b = self.array[3:5]
            self.c = self.array[5]
            self.ξ = self.array[6:]
        elif len(args) >= 1:
            assert isinstance(args[0], (int, np.ndarray))
            self.ξ = np.zeros(args[0]) if isinstance(args[0], int) else args[0].reshape(-1)
            self.a = np.array([1, 1, 0]) if len(args) < 2 else args[1].flat[np.array([0, 3, 1])]
            self.b = np.zeros(2) if len(args) < 3 else args[2].astype(float)
            self.c = 0 if len(args) < 4 else float(args[3])
            self.array = np.concatenate([self.a, self.b, np.array([self.c]), self.ξ])
        else:
            raise ValueError('Initialization failed')

    @staticmethod
    def get_model(params):
        model = params if isinstance(params, DeformableShapeModel) else DeformableShapeModel(params)
        assert not np.isnan(model.array).any()
        return model

    def copy(self):
        return DeformableShapeModel(self.array.copy())

    @property
    def A(self):
        return np.array([self.a[0], self.a[2], self.a[2], self.a[1]]).reshape((2, 2))

    def map_to_image_pixels(self, g, roi, pad=0):
        """Model in the pixel coordinates of the whole (padded) image (dsm.py:113-128)."""
        assert pad >= 0 and isinstance(pad, int)
        g_max_coord, roi_max_coord = 2 * pad + np.array(g.model.shape) - 1., np.array(roi.model.shape) - 1.
        G = np.diag(1. / roi_max_coord)
        v = -G.dot(np.add(roi.offset, pad))
        A = G.dot(self.A).dot(G)
        b = G.dot(self.A.dot(v) + self.b)
        c = np.inner(v, self.A.dot(v)) + 2 * np.inner(self.b, v) + self.c
        return DeformableShapeModel(self.ξ, A, b, c)


class _SmoothMatShape(object):
    """Stands in for ``Energy.smooth_mat`` where callers only read its shape (objects.py:380,385,393)."""

    def __init__(self, n, m): self.shape = (n, m)


class Energy(object):
    """Handle on the energy of one fitted region, as returned by :func:`superdsm_b200.objects.cvxprog`.

    ``J(model)`` is evaluated on the device (``sdsm_eval_candidate``); the value at the fitted model is cached.
    """

    def __init__(self, engine, index, n, m, solution, energy):
        self._engine, self._index = engine, index
        self.smooth_mat = _SmoothMatShape(n, m)
        self._solution = np.asarray(solution, dtype=float)
        self._energy = float(energy)

    def __call__(self, params):
        arr = DeformableShapeModel.get_model(params).array
        if arr.shape == self._solution.shape and np.array_equal(arr, self._solution): return self._energy
        return self._engine.eval_candidate(self._index, arr)[0]

    def grad(self, params):
        return self._engine.eval_candidate(self._index, DeformableShapeModel.get_model(params).array)[1]

    def hessian(self, params):
        return self._engine.eval_candidate(self._index, DeformableShapeModel.get_model(params).array)[2]
