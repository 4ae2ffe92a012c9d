"""Candidate-sharded ``compute_objects`` for one process per GPU.

The candidates of one ``compute_objects`` call are independent, so they are split over the ranks by cost (LPT), every
rank solves its shard on its own device, and the small per-candidate states are exchanged with ``all_gather_object`` --
host-side plumbing, no collective on the data path.  All ranks end up with all objects computed, which is what the
unchanged ``MinSetCover`` stage of the reference needs (``globalenergymin.py:357-364``).
"""
import numpy as np

from .engine import lpt_shards


def candidate_costs(objects, atom_pixel_count):
    """Cost model of one fit: pixels of the region (sum over the footprint's atoms), the dominant factor of both the
    evaluation work and the number of deformation parameters."""
    return np.array([float(sum(atom_pixel_count[lab] for lab in obj.footprint)) for obj in objects])


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
    local = [(int(mine[k]), state, bool(fb)) for k, state, fb in solve_shard([objects[i] for i in mine], y, atoms, dsm_cfg, log_root_dir)]
    if world > 1:
        gathered = [None] * world
        dist.all_gather_object(gathered, local, group=group)
    else:
        gathered = [local]
    fallbacks, seen = 0, 0
    for part in gathered:
        for cidx, state, fb in part:
            objects[cidx].set(state)
            fallbacks += int(fb)
            seen += 1
    if seen != len(objects): raise RuntimeError(f'sharded compute_objects: {seen} of {len(objects)} results arrived')
    return fallbacks
