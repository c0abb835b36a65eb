"""BASELINE.json configs[4]: multi-object sweep, 8 objects x 512 query frames at the headline shape, sharded over the GPUs of one box.

    python tools/bench_sweep.py                                  # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/bench_sweep.py

Objects are dealt round-robin by ``onepose_b200.shard`` (an object's DB lives on one GPU), every rank walks its objects with the
device-side sequence loop (``evaluator.run_sequence``: upload from pinned host memory, match, gather, PnP, cm-degree errors), and only
the per-item results are gathered on rank 0 (gloo side group, outside the timed region).  No collective on the data path.  Each
object has 64 distinct synthetic frames, cycled to 512 (every frame is uploaded and processed; nothing is cached between blocks
except the registered object DB).  Timed on the device between barriers, max over ranks.
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

from onepose_b200 import objectdb, shard, synth
from onepose_b200.evaluator import Evaluator, run_sequence
from onepose_b200.matcher import GATsSuperGlue

OBJECTS, FRAMES, DISTINCT, N2, N3, L = int(os.environ.get("OBJECTS", 8)), int(os.environ.get("FRAMES", 512)), 64, 2048, 8192, 8


def main():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    side = None
    if world > 1:
        sys.stdout.flush()                 # the "NCCL version ..." banner of communicator creation goes to stdout: send it to stderr
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0))))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
            os.close(saved_stdout)
        side = dist.new_group(backend="gloo")
    model = GATsSuperGlue(dict(synth.DEFAULT_HPARAMS)).cuda().eval()
    model.load_state_dict(synth.make_weights(0, damped=True), strict=True)
    # SWEEP_GRAPHS=1: CUDA-graph serving mode inside run_sequence (one graph per object DB, captured by that object's first block;
    # the matcher keeps two graphs alive, so this pays when a rank serves one or two objects -- the 4- and 8-GPU runs)
    graphs = os.environ.get("SWEEP_GRAPHS", "0") == "1"
    model.enable_graphs(graphs)
    plan = shard.assign([FRAMES] * OBJECTS, world, block=FRAMES)[rank]        # one work item per object: (object, 0, FRAMES)
    dbs, seqs = {}, {}
    for obj_id, _, _ in plan:
        obj = synth.make_object(obj_id, N3, L)
        frames = [synth.make_frame(obj, f, N2) for f in range(DISTINCT)]
        q = synth.pack_batch(obj, frames)
        dbs[obj_id] = objectdb.ObjectDB(torch.from_numpy(obj.keypoints3d).cuda(), q["descriptors3d_db"][0].t().contiguous().cuda(),
                                        q["descriptors2d_db"][0].t().contiguous().cuda(), L)
        rep = FRAMES // DISTINCT
        seqs[obj_id] = (q["keypoints2d"].repeat(rep, 1, 1).pin_memory(), q["descriptors2d_query"].repeat(rep, 1, 1).pin_memory(),
                        frames[0].K, np.tile(np.stack([f.pose_gt for f in frames]), (rep, 1, 1)))

    def process(obj_id, first, n):
        kp, ds, K, gt = seqs[obj_id]
        ev = Evaluator()
        res = run_sequence(model, dbs[obj_id], kp[first:first + n], ds[first:first + n], K, gt[first:first + n], evaluator=ev)
        return {"frames": n, "with_pose": int((res["n_inliers"] > 0).sum()), "eval": res["eval"]}

    if plan:                                                                   # warm-up: one block of the first object (with graphs: of every object this rank keeps a graph for)
        for o in ([p[0] for p in plan][:2] if graphs else [plan[0][0]]):
            run_sequence(model, dbs[o], seqs[o][0][:64], seqs[o][1][:64], seqs[o][2], seqs[o][3][:64])
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    local = {item: process(*item) for item in plan}
    e.record()
    torch.cuda.synchronize()
    ms = torch.tensor([s.elapsed_time(e)], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.barrier()
    wall = time.perf_counter() - t0
    merged = shard.gather_results(local, rank, world, group=side)
    if rank == 0:
        total = sum(v["frames"] for v in merged.values())
        print(json.dumps({
            "metric": "query frames/sec (match+PnP), multi-object sweep", "value": total / (float(ms) * 1e-3), "unit": "frames/s",
            "n_gpus": world, "cuda_graphs": graphs, "ms": float(ms), "wall_ms_rank0": wall * 1e3, "scaling": "strong",
            "config": {"workload": f"{OBJECTS} objects x {FRAMES} frames, N2d={N2}, N3d={N3} x {L} leaves, objects dealt round-robin, "
                                   "frames uploaded from pinned host memory in blocks of 64, evaluator on the device",
                       "data": f"synthetic; {DISTINCT} distinct frames per object cycled to {FRAMES}"},
            "frames_with_pose": sum(v["with_pose"] for v in merged.values()),
            "eval_per_object": {str(k[0]): v["eval"] for k, v in merged.items()},
        }))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
