"""Candidate-sharded ``compute_objects`` for one process per GPU.

The candidates of one ``compute_objects`` call are independent, so they are split over the ranks by cost (LPT), every
rank solves its shard on its own device (each holds a replica of the image), and the small per-candidate results travel as ONE
packed record per rank through ``all_gather_object`` (numpy arrays: energies, flags, offsets, bit-packed foreground
fragments -- no per-object pickling) -- host-side plumbing, no collective on the data path.  All ranks end up with all
objects computed, which is what the unchanged ``MinSetCover`` stage of the reference needs on every rank
(``globalenergymin.py:357-364``).
"""
import numpy as np

from .engine import lpt_shards


def candidate_costs_from_counts(footprints, atom_pixel_count):
    """Cost model of one fit: pixels of the region (sum over the footprint's atoms), the dominant factor of both the
    evaluation work and the number of deformation parameters."""
    return np.array([float(sum(atom_pixel_count[lab] for lab in fp)) for fp in footprints])


def candidate_costs(objects, atom_pixel_count):
    return candidate_costs_from_counts([obj.footprint for obj in objects], atom_pixel_count)


def pack_states(indices, states):
    """One record for a shard's results: ``states`` as ``objects._compute_objects`` yields them."""
    n = len(indices)
    rec = dict(index=np.asarray(indices, np.int64), energy=np.zeros(n), flags=np.zeros(n, np.uint8), time=np.zeros(n, np.float32),
               offset=np.zeros((n, 2), np.int32), shape=np.zeros((n, 2), np.int32))
    bits = []
    for k, (state, fb) in enumerate(states):
        rec['energy'][k], rec['time'][k] = state.energy, state.processing_time
        rec['flags'][k] = (1 if state.on_boundary else 0) | (2 if state.is_optimal else 0) | (4 if fb else 0)
        rec['offset'][k], rec['shape'][k] = state.fg_offset, state.fg_fragment.shape
        bits.append(np.packbits(np.asarray(state.fg_fragment, bool)))
    rec['bits'] = np.concatenate(bits) if bits else np.zeros(0, np.uint8)
    return rec


def unpack_states(rec, objects):
    """Applies a record to the caller's objects (``Object.set``); returns the number of fallbacks in it."""
    from .objects import ObjectState
    pos, fallbacks = 0, 0
    for k, cidx in enumerate(rec['index']):
        h, w = (int(v) for v in rec['shape'][k])
        nbytes = (h * w + 7) // 8
        st = ObjectState()
        st.footprint = set(objects[cidx].footprint)
        st.energy, st.processing_time = float(rec['energy'][k]), float(rec['time'][k])
        f = int(rec['flags'][k])
        st.on_boundary, st.is_optimal = bool(f & 1), bool(f & 2)
        st.fg_offset = rec['offset'][k].astype(int)
        st.fg_fragment = np.unpackbits(rec['bits'][pos:pos + nbytes])[:h * w].reshape(h, w).astype(bool)
        pos += nbytes
        objects[cidx].set(st)
        fallbacks += (f >> 2) & 1
    return fallbacks


def compute_objects_sharded(objects, y, atoms, dsm_cfg, log_root_dir=None, group=None, solve_shard=None):
    """Like ``objects.compute_objects`` but candidate-sharded over the ranks of ``group`` (default: the world group).

    ``solve_shard(shard_objects, y, atoms, dsm_cfg, log_root_dir)`` yields ``(index in shard, state, fallback)``; it defaults
    to the CUDA path (``objects._compute_objects``).  Returns the number of fallbacks over all ranks.
    """
    import torch.distributed as dist
    objects = list(objects)
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    if solve_shard is None:
        from . import objects as _objects
        def solve_shard(objs, y_, atoms_, cfg_, log_):
            return _objects._compute_objects(objs, y_, atoms_, None, None, cfg_, log_)
    count = np.bincount(np.asarray(atoms).reshape(-1))
    shards = lpt_shards(candidate_costs(objects, count), world)
    mine = shards[rank]
    local = sorted(((int(k), state, bool(fb)) for k, state, fb in solve_shard([objects[i] for i in mine], y, atoms, dsm_cfg, log_root_dir)), key=lambda t: t[0])
    rec = pack_states([int(mine[k]) for k, _, _ in local], [(state, fb) for _, state, fb in local])
    if world > 1:
        gathered = [None] * world
        dist.all_gather_object(gathered, rec, group=group)
    else:
        gathered = [rec]
    fallbacks = sum(unpack_states(part, objects) for part in gathered)
    seen = sum(len(part['index']) for part in gathered)
    if seen != len(objects): raise RuntimeError(f'sharded compute_objects: {seen} of {len(objects)} results arrived')
    return fallbacks
