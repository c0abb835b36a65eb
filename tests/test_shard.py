"""Multi-process (gloo, world_size 2, CPU) tests of the data-parallel sharding logic."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from onepose_b200 import shard


def _fake_process(obj, first, n):
    # stands in for match + PnP: a deterministic per-frame "pose" that depends only on (object, frame)
    f = np.arange(first, first + n, dtype=np.float64)
    return {"pose": np.stack([np.full((3, 4), obj * 1000.0) + fi for fi in f]), "frames": f.astype(np.int64)}


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, frames_per_object, block, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    res = shard.run_sharded(_fake_process, frames_per_object, rank, world, block)
    if rank == 0:
        torch.save(res, out)
    else:
        assert res is None
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("frames_per_object", [[512] * 8, [100], [5, 0, 130]])
@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_assignment_partitions_the_work(frames_per_object, world):
    per_rank = shard.assign(frames_per_object, world, 64)
    flat = [it for r in per_rank for it in r]
    assert sorted(flat) == sorted(shard.make_work_items(frames_per_object, 64))      # every item exactly once
    assert len(set(flat)) == len(flat)
    if len(frames_per_object) >= world:                                               # an object lives on one GPU
        for obj in range(len(frames_per_object)):
            assert len({r for r, items in enumerate(per_rank) for it in items if it[0] == obj}) <= 1
    covered = sum(n for _, _, n in flat)
    assert covered == sum(frames_per_object)


@pytest.mark.parametrize("frames_per_object", [[70, 64, 3], [200]])
def test_two_ranks_return_the_single_rank_result(tmp_path, frames_per_object):
    single = shard.run_sharded(_fake_process, frames_per_object, 0, 1, 64)
    out = str(tmp_path / "merged.pt")
    mp.spawn(_worker, args=(2, _free_port(), frames_per_object, 64, out), nprocs=2, join=True)
    merged = torch.load(out, weights_only=False)
    assert list(merged) == list(single)
    for k in single:
        assert np.array_equal(merged[k]["pose"], single[k]["pose"]) and np.array_equal(merged[k]["frames"], single[k]["frames"])
