"""Loads the reference's own ``superdsm/dsm.py``, ``image.py``, ``_aux.py`` and ``objects.py`` UNCHANGED.

TEST INFRASTRUCTURE, and only usable where ``/root/reference`` exists (the build container, never the GPU box).
It is the recipe of SURVEY.md Appendix A: the package ``__init__`` is bypassed (it imports the whole pipeline,
``superdsm/__init__.py:9-10``) and the absent third-party modules are stubbed:

* ``ray``: serial identity stubs (``objects.py:215``),
* ``cvxopt``: ``matrix`` / ``spmatrix`` / ``solvers.cp`` bound to ``oracle/cpsolver.py`` (the restated solver;
  cvxopt itself is not installable here),
* ``skimage.util.view_as_windows`` = ``numpy.lib.stride_tricks.sliding_window_view`` (``dsm.py:156``),
* ``skimage.morphology``: imported by ``objects.py:9`` but not used on this path,
* ``IPython.display.clear_output``: no-op (``output.py:6``),
* ``superdsm._mkl``: ``dot = A @ B``, ``gram = triu(A @ A.T)`` (MKL's syrk returns the upper triangle,
  ``_libs/sparse_dot_mkl/_sparse_sparse.py:37-57``); libmkl_rt is not installed here.

Nothing is copied from the reference: its files are executed where they lie.
"""
import importlib.util
import os
import sys
import types
import warnings

import numpy as np
import scipy.sparse

REFERENCE_ROOT = os.environ.get('SUPERDSM_REFERENCE_ROOT', '/root/reference')


def available():
    return os.path.isfile(os.path.join(REFERENCE_ROOT, 'superdsm', 'dsm.py'))


def _stub(name, **attrs):
    mod = types.ModuleType(name)
    for key, value in attrs.items(): setattr(mod, key, value)
    sys.modules[name] = mod
    return mod


def load():
    """Returns a namespace with the reference modules ``dsm``, ``image``, ``objects``, ``_aux``."""
    if 'superdsm.dsm' in sys.modules and getattr(sys.modules['superdsm'], '_oracle_harness', False):
        pkg = sys.modules['superdsm']
        return types.SimpleNamespace(dsm=pkg.dsm, image=pkg.image, objects=pkg.objects, _aux=pkg._aux)
    assert available(), f'reference not found under {REFERENCE_ROOT}'
    from . import cpsolver

    class _Remote(object):
        def __init__(self, f): self.f = f
        def __call__(self, *a, **kw): return self.f(*a, **kw)
        def remote(self, *a, **kw): return self.f(*a, **kw)

    _stub('ray', remote=lambda f: _Remote(f), put=lambda x: x, get=lambda x: x,
          wait=lambda ids, num_returns=1: (ids[:num_returns], ids[num_returns:]))
    solvers = _stub('cvxopt.solvers', cp=cpsolver.cp_cvxopt_protocol, options={})
    _stub('cvxopt', matrix=cpsolver.matrix, spmatrix=cpsolver.spmatrix, solvers=solvers)
    util = _stub('skimage.util', view_as_windows=lambda arr, shape: np.lib.stride_tricks.sliding_window_view(arr, shape))
    morph = _stub('skimage.morphology')
    _stub('skimage', util=util, morphology=morph)
    display = _stub('IPython.display', clear_output=lambda *a, **kw: None)
    _stub('IPython', display=display)

    pkg = types.ModuleType('superdsm')
    pkg.__path__ = [os.path.join(REFERENCE_ROOT, 'superdsm')]
    pkg._oracle_harness = True
    sys.modules['superdsm'] = pkg
    _stub('superdsm._mkl', dot=lambda A, B: A @ B, gram=lambda A: scipy.sparse.triu(A @ A.T, format='csr'))
    _stub('superdsm.io', imread=None)

    def load_file(name):
        path = os.path.join(REFERENCE_ROOT, 'superdsm', name + '.py')
        spec = importlib.util.spec_from_file_location('superdsm.' + name, path)
        mod = importlib.util.module_from_spec(spec)
        sys.modules['superdsm.' + name] = mod
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            spec.loader.exec_module(mod)
        setattr(pkg, name, mod)
        return mod

    aux = load_file('_aux')
    load_file('output')
    dsm = load_file('dsm')
    image = load_file('image')
    objects = load_file('objects')
    objects._compute_objects._DEBUG = True      # serial path, no Ray (objects.py:271-273)
    return types.SimpleNamespace(dsm=dsm, image=image, objects=objects, _aux=aux)
