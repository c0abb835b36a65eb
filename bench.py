#!/usr/bin/env python
"""Benchmark of the GATsSPG match + PnP hot path (BASELINE.json metric: query frames/sec).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Own arm: a "step" is one pass of the hot path over one batch of 64 synthetic query frames of one object
at the headline shape (N2d=2048 query keypoints, N3d=8192 3D points x 8 leaves): GATsSuperGlue.forward
-> valid-match gather -> batched PnP-RANSAC, all through the C ABI of libonepose_b200.so.  ``value`` is
frames/s with inputs resident in HBM; ``e2e`` is the same step fed from pinned HOST buffers through the
drop-in module (query descriptors/keypoints copied host->device and matches/poses copied back every
step; the static object DB stays resident, as a caller registering an object once would keep it).
N > 1: one process per GPU (torchrun), every rank runs its own object with the same per-GPU batch (weak
scaling, no collective on the data path); time = max over ranks of CUDA-event time.

Reference arm (--impl reference): the reference is pure Python and is not present on the GPU box, so the arm
times the CPU restatement (oracle/: torch-CPU matcher with the reference's op sequence + cv2.solvePnPRansac,
the very library call the reference makes) on the host cores, one frame (B=1, the reference's own usage,
inference.py:146) per step.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "query frames/sec (match+PnP) at N2d=2048,N3d=8192"
UNIT = "frames/s"
N2, N3, LEAVES, FRAMES = 2048, 8192, 8, 64
PNP_HYPOTHESES = 4096


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return dict(hbm=p["hbm_gbs"], tensor_burst=p["bf16_tflops"], tensor=p.get("bf16_tflops_sustained", p["bf16_tflops"]),
                    source="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, tensor_burst=1590.0, tensor=1400.0, source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 25 ms from the last warm-up steps to the end of the timed region (the sampler
    process needs ~0.1 s to come up, as long as a default timed region)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.lines, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "25"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=lambda: [self.lines.append(l) for l in self.proc.stdout], daemon=True).start()
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        time.sleep(0.05)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------- CPU side
def cpu_frame_runner(object_id=0):
    """Returns (run_one_frame(f) -> n_matches, cores): the oracle matcher + the reference wrapper over cv2."""
    import torch
    from onepose_b200 import synth
    from oracle import gats_spg_oracle as orc
    from oracle import pnp_oracle as po

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    hp = dict(synth.DEFAULT_HPARAMS)
    w = synth.make_weights(0, damped=True)
    obj = synth.make_object(object_id, N3, LEAVES)

    def run(f):
        fr = synth.make_frame(obj, f, N2)
        data = synth.pack_batch(obj, [fr])
        with torch.no_grad():
            pred, _ = orc.matcher_forward(data, w, hp)
        mk2, mk3, _ = orc.gather_matches(pred["matches0"].numpy(), pred["matching_scores0"].numpy(),
                                         fr.keypoints2d, obj.keypoints3d)
        po.ransac_pnp_cv2(fr.K, mk2, mk3, scale=1000)
        return len(mk2)

    return run, cores


def time_cpu(steps, warmup):
    run, cores = cpu_frame_runner()
    for f in range(warmup):
        run(f)
    t0 = time.perf_counter()
    for f in range(steps):
        run(1000 + f)
    dt = time.perf_counter() - t0
    return dt, cores


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return                                      # rank 0 alone runs the CPU arm
    dt, cores = time_cpu(args.steps, args.warmup)
    fps = args.steps / dt
    sample = f"{args.steps} single-frame steps (B=1) of the headline workload after {args.warmup} warm-up"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": fps, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"headline: N2d={N2}, N3d={N3} x {LEAVES} leaves, 1 frame per step (B=1), match + gather + "
                               "cv2.solvePnPRansac(EPNP, 5 px, 10000 it)", "weights": "random-init, damped (SURVEY.md 8c)"},
        "cpu_baseline": {"value": fps, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": fps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# ---------------------------------------------------------------------------------------------- own arm
def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from onepose_b200 import synth
    from onepose_b200._lib import timeline
    from onepose_b200.matcher import GATsSuperGlue

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # the "NCCL version ..." banner of communicator creation goes to stdout, in front of the JSON line: send it to stderr
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
            os.close(saved_stdout)
    peaks = load_peaks()

    # --- synthetic workload: this rank's object + FRAMES query frames (seeded) ---
    hp = dict(synth.DEFAULT_HPARAMS)
    model = GATsSuperGlue(hp).to(dev).eval()
    model.load_state_dict(synth.make_weights(0, damped=True), strict=True)
    obj = synth.make_object(rank, N3, LEAVES)
    frames = [synth.make_frame(obj, f, N2) for f in range(FRAMES)]
    host = synth.pack_batch(obj, frames)                       # CPU tensors, DB expand()-ed over the batch
    K_dev = torch.from_numpy(frames[0].K)[None].to(dev)
    db_dev = {k: host[k][:1].to(dev).expand(FRAMES, *host[k].shape[1:]) for k in
              ("keypoints3d", "descriptors3d_db", "descriptors2d_db")}
    q_pin = {k: host[k].contiguous().pin_memory() for k in ("keypoints2d", "descriptors2d_query")}
    resident = dict(db_dev, **{k: v.to(dev) for k, v in q_pin.items()}, query_intrinsic=K_dev)
    out_pin = dict(pose=torch.empty(FRAMES, 3, 4, dtype=torch.float64).pin_memory(),
                   n_inliers=torch.empty(FRAMES, dtype=torch.int32).pin_memory(),
                   matches0=torch.empty(FRAMES, N2, dtype=torch.int64).pin_memory(),
                   scores0=torch.empty(FRAMES, N2, dtype=torch.float32).pin_memory())
    h2d_bytes = sum(v.numel() * v.element_size() for v in q_pin.values())
    d2h_bytes = sum(v.numel() * v.element_size() for v in out_pin.values())

    def step_resident():
        pred, conf = model(resident)
        return pred

    # End to end: every step uploads its 64 query frames from pinned host memory and reads its results back.  The upload of
    # step i+1 is issued on a copy stream while step i computes (a double-buffered input pipeline, what a frame loader
    # does); the first upload of a timed region is not hidden, every upload is inside the region.
    copy_stream = torch.cuda.Stream(device=dev)
    pending = []

    def upload():
        with torch.cuda.stream(copy_stream):
            up = {k: v.to(dev, non_blocking=True) for k, v in q_pin.items()}
            pending.append((up, copy_stream.record_event()))

    def step_e2e(prefetch_next=True):
        if not pending:
            upload()
        up, ready = pending.pop(0)
        main = torch.cuda.current_stream(dev)
        main.wait_event(ready)
        for v in up.values():
            v.record_stream(main)
        if prefetch_next:
            upload()
        data = dict(db_dev, query_intrinsic=K_dev, **up)
        pred, conf = model(data)
        with torch.cuda.stream(model.pose_stream(dev) if model.pnp_overlap else main):     # the pose is read back behind its PnP
            out_pin["pose"].copy_(pred["pose"], non_blocking=True)
            out_pin["n_inliers"].copy_(pred["n_inliers"], non_blocking=True)
        out_pin["matches0"].copy_(pred["matches0_batched"], non_blocking=True)
        out_pin["scores0"].copy_(pred["matching_scores0_batched"], non_blocking=True)
        return pred

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, instrument=False):
        barrier()
        timeline.reset()
        timeline.enabled = instrument
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(steps):
            pred = fn()
        e.record()
        barrier()
        timeline.enabled = False
        ms = torch.tensor([s.elapsed_time(e)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms.item(), pred

    # Steps are pipelined as a frame server would run them: gather + PnP of step i go to the module's pose stream and overlap the
    # matcher of step i+1 (GATsSuperGlue.pnp_overlap); every step's work is complete at the barrier that ends the timed region.
    model.pnp_overlap = not args.no_pnp_overlap
    clocks = ClockSampler(local_rank)
    clocks.start()
    for _ in range(max(args.warmup, 3)):
        step_resident()
    ms, pred = timed(step_resident, args.steps)
    clk = clocks.stop()
    # per-kernel durations (roofline, shares, launch count): the same K steps once more with CUDA events around every
    # launch and the 2D / 3D streams serialised on one CUDA stream, so that no launch's interval contains another's
    model.two_streams = False
    overlap = model.pnp_overlap
    model.pnp_overlap = False
    step_resident()
    ms_serial, _ = timed(step_resident, args.steps, instrument=True)
    model.two_streams = True
    model.pnp_overlap = overlap
    summary = timeline.summary()
    launches = timeline.launches // args.steps
    for _ in range(2):
        step_e2e(prefetch_next=False)
    e2e_calls = [0]

    def e2e_step_counted():                     # exactly K uploads inside the timed region: the last step prefetches nothing
        e2e_calls[0] += 1
        return step_e2e(prefetch_next=e2e_calls[0] < args.steps)

    ms_e2e, pred_e = timed(e2e_step_counted, args.steps)

    total_frames = FRAMES * args.steps * world
    value = total_frames / (ms / 1e3)
    e2e_value = total_frames / (ms_e2e / 1e3)
    n_inl = pred["n_inliers"].float()
    sanity = {"frames_with_pose": int((pred["n_inliers"] > 0).sum()), "mean_matches": float(pred["n_matches"].float().mean()),
              "mean_inliers": float(n_inl.mean())}

    if rank == 0:
        # --- per-kernel time shares + roofline of the dominant kernel group -----------------------
        tot = sum(v[0] for v in summary.values())
        kernels = {}
        for name, (kms, n, work) in sorted(summary.items(), key=lambda kv: -kv[1][0]):
            ent = {"ms_per_step": kms / args.steps, "calls_per_step": n / args.steps, "share": kms / tot}
            if work > 0:
                if name.startswith(("op_linear", "op_atb")):
                    ent.update(bound="tensor", achieved=work / (kms * 1e-3) / 1e12, peak=peaks["tensor"], unit="TFLOP/s")
                else:
                    ent.update(bound="hbm", achieved=work / (kms * 1e-3) / 1e9, peak=peaks["hbm"], unit="GB/s")
                ent["frac"] = ent["achieved"] / ent["peak"]
            kernels[name] = ent
        dom = next(k for k in kernels if "frac" in kernels[k])
        d = kernels[dom]
        try:        # DRAM bytes per launch of that kernel from the committed ncu capture (profiles/traffic.json), else null
            with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "traffic.json")) as f:
                traffic = json.load(f)
        except (OSError, ValueError):
            traffic = {}
        tr = traffic.get(dom)
        roofline = {"kernel": dom, "bound": d["bound"], "achieved": d["achieved"], "peak": d["peak"], "unit": d["unit"],
                    "frac": d["frac"], "traffic": None if tr is None else tr["traffic_bytes"],
                    "traffic_launch": None if tr is None else tr["launch"], "peak_source": peaks["source"] + (", sustained" if d["bound"] == "tensor" else ""),
                    "share_of_step": d["share"]}
        if d["bound"] == "tensor":
            # fp32-class accuracy from fp16 MMAs costs three MMAs per algorithmic product (hi*hi + lo*hi + hi*lo): the tensor pipe
            # executes 3x the algorithmic flops that `achieved` counts
            roofline.update(mma_passes=3, achieved_issued=3 * d["achieved"], frac_issued=3 * d["frac"])
        sim = kernels.get("op_sim_dual_softmax_planes") or kernels.get("op_sim_dual_softmax_rows") or kernels.get("op_sim_dual_softmax")
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            n_cpu = 4
            dt, cores = time_cpu(n_cpu, 1)
            cpu = {"value": n_cpu / dt, "unit": UNIT, "cores": cores, "kind": "port",
                   "sample": f"{n_cpu} frames (B=1 per call) of the headline workload after 1 warm-up: oracle torch-CPU matcher "
                             "+ cv2.solvePnPRansac"}
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"headline: {FRAMES} frames/step of one object per GPU, N2d={N2}, N3d={N3} x {LEAVES} leaves, "
                                   f"match + gather + PnP-RANSAC ({PNP_HYPOTHESES} hypotheses/frame)",
                       "weights": "random-init, damped (SURVEY.md 8c) so PnP sees real correspondences",
                       "l2": "working set per step (~6 GB of activations + 4 GiB conf) exceeds the 126 MB L2; no flush needed",
                       "pipeline": "gather + PnP of step i on a pose stream, overlapping the matcher of step i+1" if model.pnp_overlap else "none",
                       "db_resident": True, "parallelism": f"dp{world} (objects sharded, no collective)"},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes,
                    "ms_per_step": ms_e2e / args.steps,
                    "h2d": "pinned host -> device on a copy stream, step i+1 uploaded while step i computes (all K uploads in the timed region)"},
            "gpu_launches": launches * args.steps, "gpu_launches_per_step": launches,
            "kernel_timing": {"how": "CUDA events around every launch in a second pass of the same steps with the two matcher "
                                     "streams serialised on one CUDA stream", "ms_per_step": ms_serial / args.steps},
            "roofline": roofline,
            "roofline_sim_dual_softmax": None if sim is None else dict(
                {k: sim[k] for k in ("bound", "achieved", "peak", "unit", "frac", "ms_per_step", "share")},
                traffic=(traffic.get("op_sim_dual_softmax_planes") or {}).get("traffic_bytes"),
                # the similarity is computed twice (statistics pass, write pass); a product is one fp16 MMA plus two fp8 MMAs at twice
                # the rate (= 2 fp16-equivalent passes; 3 with OPB_SIM_MMA=f16 / OPB_SIM_CG=1)
                mma_passes_fp16_equivalent=SIM_MMA_UNITS,
                tensor_issued_tflops=2 * SIM_MMA_UNITS * 2.0 * FRAMES * N2 * N3 * 256 / (sim["ms_per_step"] * 1e-3) / 1e12,
                tensor_frac_issued=2 * SIM_MMA_UNITS * 2.0 * FRAMES * N2 * N3 * 256 / (sim["ms_per_step"] * 1e-3) / 1e12 / peaks["tensor"]),
            "cpu_baseline": cpu, "clocks": clk, "kernels": kernels, "sanity": sanity,
        }))
    if world > 1:
        dist.destroy_process_group()


SIM_MMA_UNITS = 3 if (os.environ.get("OPB_SIM_MMA") == "f16" or os.environ.get("OPB_SIM_CG") == "1") else 2


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--no-pnp-overlap", action="store_true", help="gather + PnP on the matcher's stream (no pipelining across steps)")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 and world == 1:
        # convenience: python bench.py --gpus N re-launches itself under torchrun
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    run_ours(args)


if __name__ == "__main__":
    main()
