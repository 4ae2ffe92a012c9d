"""What ``objects.cvxprog`` hands back to its callers: the fitted parameters and a handle on the energy.

The arithmetic of the reference's ``dsm.Energy`` / ``dsm.CP`` lives in the CUDA library.  Callers of the seam only read the
parameter vector (``model.array``, layout of ``dsm.py:48-63``: a1, a2, a3, b1, b2, c, xi...) and evaluate ``J(model)``
(``c2freganal.py:74``, ``objects.py:399``), so that is all these two classes offer.
"""
import numpy as np


def _as_array(params):
    arr = getattr(params, 'array', params)
    arr = np.asarray(arr, dtype=float).reshape(-1)
    if len(arr) < 6 or np.isnan(arr).any(): raise ValueError('a model needs at least six finite parameters')
    return arr


class ModelParams(object):
    """Parameter vector of a fitted model with the reference's field names as read-only views."""

    def __init__(self, array):
        self.array = _as_array(array).copy()

    a = property(lambda self: self.array[:3])
    b = property(lambda self: self.array[3:5])
    c = property(lambda self: self.array[5])
    ξ = property(lambda self: self.array[6:])

    @property
    def A(self):
        a1, a2, a3 = self.array[:3]
        return np.array([[a1, a3], [a3, a2]])

    def copy(self):
        return ModelParams(self.array)


class _Shape(object):
    def __init__(self, n, m): self.shape = (n, m)


class Energy(object):
    """``J`` of one fitted region.  ``J(model)``, ``J.grad(model)``, ``J.hessian(model)`` are evaluated on the device
    (``sdsm_eval_candidate``) while the batch the region was solved in is the engine's last one; the value at the fitted
    model is known without a call.  ``J.smooth_mat.shape`` is (n, m) as callers read it (``objects.py:380``)."""

    def __init__(self, engine, index, n, m, solution, energy):
        self._engine, self._index, self._token = engine, index, engine._last
        self.smooth_mat = _Shape(n, m)
        self._solution = np.asarray(solution, dtype=float)
        self._energy = float(energy)

    def _eval(self, params, hessian):
        if self._engine._last is not self._token:
            raise RuntimeError('this Energy handle belongs to a batch that is no longer resident on the device')
        return self._engine.eval_candidate(self._index, _as_array(params), hessian=hessian)

    def __call__(self, params):
        arr = _as_array(params)
        if arr.shape == self._solution.shape and np.array_equal(arr, self._solution): return self._energy
        return self._eval(arr, False)[0]

    def grad(self, params): return self._eval(params, False)[1]

    def hessian(self, params): return self._eval(params, True)[2]
