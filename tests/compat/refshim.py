"""Caller-side stand-ins for the GPU box, where ``/root/reference`` does not exist: the two classes a caller of the seam
hands in.  TEST INFRASTRUCTURE -- the product (``superdsm_b200/``) is duck-typed and never imports this; in the build
container ``tests/test_reference_seam.py`` runs the same seam with the reference's own ``Image`` / ``Object`` classes.

Only what the tests need: an image is (model, mask[, offset, full_mask]); an object is a footprint plus result fields.
"""
import numpy as np
import scipy.ndimage as ndi


class Image(object):
    def __init__(self, model, mask=None, full_mask=None, offset=(0, 0)):
        self.model = model
        self.mask = np.ones(model.shape, bool) if mask is None else mask
        self.full_mask = self.mask if full_mask is None else full_mask
        self.offset = tuple(offset)

    def window(self):
        """Slices of this (cropped) image inside its parent."""
        (r, c), (h, w) = self.offset, self.mask.shape
        return np.s_[r:r + h, c:c + w]

    def shrink_mask(self, mask):
        return mask[self.window()]

    def get_region(self, mask, shrink=False):
        """Sub-image restricted to `mask` (semantics of the reference's ``Image.get_region``, image.py:75-87)."""
        mask = self.mask & mask
        if not shrink: return Image(self.model, mask)
        rows, cols = np.flatnonzero(mask.any(axis=1)), np.flatnonzero(mask.any(axis=0))
        sel = np.s_[rows[0]:rows[-1] + 1, cols[0]:cols[-1] + 1]
        return Image(self.model[sel], mask[sel], full_mask=mask, offset=(rows[0], cols[0]))


class Object(object):
    FIELDS = ('energy', 'on_boundary', 'is_optimal', 'processing_time')

    def __init__(self, footprint=()):
        self.footprint = set(footprint)
        self.fg_offset = self.fg_fragment = None
        for f in self.FIELDS: setattr(self, f, np.nan)

    def set(self, state):
        self.footprint = set(state.footprint)
        self.fg_offset = None if state.fg_offset is None else np.array(state.fg_offset)
        self.fg_fragment = None if state.fg_fragment is None else np.array(state.fg_fragment)
        for f in self.FIELDS: setattr(self, f, getattr(state, f))
        return self

    def get_mask(self, atoms):
        return np.isin(atoms, sorted(self.footprint))

    def get_cvxprog_region(self, y, atoms, background_margin):
        """objects.py:95-128: the object's region with the pixels far from any foreground removed."""
        region = y.get_region(self.get_mask(atoms))
        region.mask = region.mask & (ndi.distance_transform_edt(y.model <= 0) <= background_margin)
        return region

    def fill_foreground(self, out, value=True):
        r, c = self.fg_offset
        h, w = self.fg_fragment.shape
        out[r:r + h, c:c + w] = value * self.fg_fragment
