"""Data-parallel sharding of the query workload over the GPUs of one box.

The reference has no multi-GPU inference path: ``inference.py:185-198`` walks (object, sequence) pairs in one
process.  Every (object, frame) is independent end to end -- InstanceNorm statistics, the linear-attention sums,
both softmax sums and RANSAC are per frame (GATs_SuperGlue.py:96-98,144-146,253; eval_utils.py:28) -- so the path
shards with NO collective on the data path: one process per GPU, a static assignment of work items, and a
host-side gather of the small per-frame results (pose 3x4 float64, inlier / match counts).  ``conf_matrix`` (64 MiB
per frame) never leaves its GPU.

Work item = (object id, first frame, number of frames), at most ``block`` frames.  Objects are dealt round-robin so
that an object's descriptor DB (72 MiB fp32 at the headline shape) is uploaded to exactly one GPU; when there are
fewer objects than ranks an object's frame blocks are dealt round-robin instead and its DB is replicated.
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

WorkItem = Tuple[int, int, int]          # (object id, first frame, n frames)


def make_work_items(frames_per_object: Sequence[int], block: int = 64) -> List[WorkItem]:
    items = []
    for obj, n in enumerate(frames_per_object):
        for start in range(0, n, block):
            items.append((obj, start, min(block, n - start)))
    return items


def assign(frames_per_object: Sequence[int], world: int, block: int = 64) -> List[List[WorkItem]]:
    """Static assignment rank -> work items (deterministic, no communication needed to agree on it)."""
    if world < 1:
        raise ValueError("world must be >= 1")
    n_obj = len(frames_per_object)
    per_rank: List[List[WorkItem]] = [[] for _ in range(world)]
    if n_obj >= world:
        for obj, n in enumerate(frames_per_object):
            for start in range(0, n, block):
                per_rank[obj % world].append((obj, start, min(block, n - start)))
    else:
        for i, item in enumerate(make_work_items(frames_per_object, block)):
            per_rank[i % world].append(item)
    return per_rank


def gather_results(local: Dict[WorkItem, dict], rank: int, world: int, group=None, dst: int = 0):
    """Host-side gather of the per-item results on ``dst`` (small tensors / numpy arrays only).

    Uses ``torch.distributed.gather_object`` on the given group (gloo on CPU; under NCCL pass a gloo side group, as the
    reference's own ``comm.gather`` does, src/utils/comm.py:177-215).  Returns the merged dict on ``dst`` (sorted by work
    item), ``None`` elsewhere.  Not part of the timed data path.
    """
    if world == 1:
        return dict(sorted(local.items()))
    import torch.distributed as dist
    buf = [None] * world if rank == dst else None
    dist.gather_object(local, buf, dst=dst, group=group)
    if rank != dst:
        return None
    merged: Dict[WorkItem, dict] = {}
    for part in buf:
        for k, v in part.items():
            if k in merged:
                raise RuntimeError(f"work item {k} was processed by two ranks")
            merged[k] = v
    return dict(sorted(merged.items()))


def run_sharded(process_item, frames_per_object: Sequence[int], rank: int, world: int, block: int = 64, group=None):
    """Runs ``process_item(obj, first, n) -> dict`` over this rank's items and gathers the results on rank 0."""
    mine = assign(frames_per_object, world, block)[rank]
    local = {item: process_item(*item) for item in mine}
    return gather_results(local, rank, world, group=group)
