"""The ``dsm/*`` hyper-parameters, unchanged from the reference (``superdsm/dsmcfg.py:6-21``)."""
import numpy as np

DSM_CONFIG_DEFAULTS = {
    'cachesize': 1,
    'cachetest': None,
    'sparsity_tol': 0,
    'init': 'elliptical',
    'smooth_amount': 10,
    'epsilon': 1.0,
    'alpha': 0.5,
    'scale': 1000,
    'smooth_subsample': 20,
    'gaussian_shape_multiplier': 2,
    'smooth_mat_dtype': 'float32',
    'smooth_mat_max_allocations': np.inf,
    'background_margin': 20,
    'cp_timeout': 300,
}


def get_dsm_cfg(cfg=None):
    """``DSM_Config.process`` (``dsmcfg.py:82-89``): the dictionary handed to the stages as ``dsm_cfg``."""
    cfg = cfg or {}
    return {key: cfg.get(key, DSM_CONFIG_DEFAULTS[key]) for key in DSM_CONFIG_DEFAULTS}
