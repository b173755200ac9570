#!/usr/bin/env python
"""bench.py — BASELINE.json's metric: 1080p SURF detect+describe frames/s (surfFast, synthetic frames).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--batch B]

A step = one pass of the hot path (integral -> Fast-Hessian -> NMS/refine -> orientation -> SURF-64) over one batch of
1080p GrayF32 frames (BASELINE.json configs[1]: 1024 frames on one B200).  N>1 (torchrun): every rank runs its own batch on
its own GPU — frames are independent, there is no collective on the data path (weak scaling); ranks only meet at the
timing barrier.

  value      frames/s with the batch already resident in HBM (CUDA events, max over ranks)
  e2e        frames/s with HOST (pinned) frames in and HOST feature lists out, copies inside the timed region
  roofline   the Fast-Hessian kernel (dominant): algorithmic bytes per launch / its CUDA-event time, vs measured HBM peak
  cpu_baseline  the CPU oracle (C++ restatement of the reference Java; no JVM on this box) on a bounded sample
--impl reference times that CPU path alone, with all host threads, and prints the same JSON shape.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H = 1920, 1080
METRIC = "1080p SURF detect+describe frames/s"
UNIT = "frames/s"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index; self.proc = None; self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=lambda: [self.lines.append(l) for l in self.proc.stdout], daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_surf_fps(frames, nthreads, min_seconds=10.0, max_frames=64):
    """frames/s of the oracle's surfFast on host cores.  frames: list of float32 [H,W] arrays, cycled."""
    from oracle import bsoracle as orc
    orc.build()
    orc.surf_fast(frames[0][:240, :320].copy(), nthreads=nthreads)      # load the .so, warm the thread pool
    n, t0 = 0, time.perf_counter()
    feats = 0
    while True:
        r = orc.surf_fast(frames[n % len(frames)], nthreads=nthreads)
        feats += len(r["x"]); n += 1
        dt = time.perf_counter() - t0
        if (dt >= min_seconds and n >= 3) or n >= max_frames:
            break
    return n / dt, n, feats / n


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU path (oracle port; the Java cannot run here) on this box's host cores."""
    if rank != 0:
        return
    from boofcv_b200.synth import blob_image
    cores = os.cpu_count() or 1
    frames = [blob_image(W, H, 234 + i) for i in range(4)]
    per_step = []
    sample_frames = 0
    for s in range(args.warmup + args.steps):
        fps, n, feats = cpu_surf_fps(frames, cores, min_seconds=4.0, max_frames=1024)
        if s >= args.warmup:
            per_step.append(fps); sample_frames += n
    v = float(np.mean(per_step))
    sample = "%d x 1080p frames per step (cycled over 4 synthetic frames), surfFast, %d threads" % (sample_frames // max(len(per_step), 1), cores)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1000.0 * (sample_frames / max(len(per_step), 1)) / v, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "surfFast detect+describe, 1920x1080 GrayF32 (BASELINE configs[1]), bounded CPU sample"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def run_gpu(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from boofcv_b200 import FactoryDetectDescribe, _native
    from boofcv_b200.synth import blob_images_torch, blob_image
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    B = args.batch
    # every rank gets its own batch (weak scaling): frame f of rank r has seed 234 + r*B + f
    frames = blob_images_torch(W, H, B, seed=234 + rank * B, device=dev)
    det = FactoryDetectDescribe.surfFast(None, None, None, np.float32, device=local_rank)
    if args.chunk:
        det.setWorkspace(args.chunk, 0)
    det.setWays(args.ways)

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()

    def step():
        det.detectBatch(frames)

    for _ in range(args.warmup):
        step()
    total_features = det.totalFeatures()          # synchronises; also proves the pool did not overflow
    launches0 = _native.lib().bs_launch_count()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    launches = _native.lib().bs_launch_count() - launches0
    # Per-kernel durations for the roofline: the timed region above spreads the batch over `ways` streams, so a stage's
    # event pair there would also count the kernels of the other streams that share the SMs with it.  The same K steps are
    # therefore run once more with every kernel on ONE stream and an event pair around each stage launch.
    timing, ms_single = None, None
    if rank == 0:
        det.setWays(1)
        step()
        det.totalFeatures()
        det.setTiming(True)
        torch.cuda.synchronize(dev)
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s0.record()
        for _ in range(args.steps):
            step()
        s1.record()
        torch.cuda.synchronize(dev)
        ms_single = s0.elapsed_time(s1)
        timing = det.getTiming()
        det.setTiming(False)
        det.setWays(args.ways)
    barrier()
    return dict(torch=torch, dist=dist, dev=dev, det=det, frames=frames, ms=ms, clocks=clocks, launches=launches,
                timing=timing, ms_single=ms_single, total_features=total_features, B=B)


def det_total_of_first(g, B):
    """features of the first B frames of the device batch (the device path's own answer for them)"""
    det = g["det"]
    det.detectBatch(g["frames"][:B])
    return det.totalFeatures()


def run_e2e(g, steps, world=1):
    """Same metric through the host-facing call (bs_surf_detect_host_batch): the step's frames live in pinned HOST memory,
    the feature lists (frame offsets, xy, scale, angle, desc, white) land in pinned HOST memory; the library overlaps the
    copies of one 32-frame sub-batch with the kernels of the other.  Returns (frames/s, h2d, d2h bytes per step)."""
    torch, det, dev = g["torch"], g["det"], g["dev"]
    from boofcv_b200 import HostFeatureBuffers
    # host memory is bounded when several ranks share one box: 1024 frames are 8.5 GB of pinned input + 6 GB of pinned output per
    # rank, so the ranks of a box share 2048 frames between them
    B = min(g["B"], max(2048 // world, 64))
    host = torch.empty((B, H, W), dtype=torch.float32, pin_memory=True)      # the step's input lives on the host
    host.copy_(g["frames"][:B])
    cap = int(g["total_features"] * 1.1 * B / g["B"]) + 4096
    out = HostFeatureBuffers(B, cap, 64, pinned=True)
    expect = int(det_total_of_first(g, B))

    def one_step():
        det.detectHostBatch(host, out)
        return B * H * W * 4, out.total * (64 * 8 + 4 * 8 + 1) + (B + 1) * 4

    one_step()                                                  # warm: worker handles, workspaces
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    for _ in range(steps):
        h2d, d2h = one_step()                                   # synchronous: returns when the results are on the host
    dt = time.perf_counter() - t0
    if out.total != expect:
        raise RuntimeError("host path found %d features, device path %d" % (out.total, expect))
    out.close()
    return B * steps / dt, h2d, d2h


def run_association(torch, dev, n=100000, reps=3):
    """Second headline of the path (BASELINE north_star): greedy association of n x n SURF-64 descriptors with backward
    validation, descriptors resident in HBM.  Tensor-core filter + exact fp64 re-score (bs_assoc mode 1)."""
    import numpy as np
    import boofcv_b200 as bs
    g = torch.Generator(device=dev); g.manual_seed(99)
    dst = torch.randn((n, 64), generator=g, device=dev, dtype=torch.float64); dst /= dst.norm(dim=1, keepdim=True)
    perm = torch.randperm(n, generator=g, device=dev)
    src = dst[perm] + 0.02 * torch.randn((n, 64), generator=g, device=dev, dtype=torch.float64)
    src /= src.norm(dim=1, keepdim=True)                       # SURF descriptors are unit vectors (UtilFeature.normalizeL2)
    a = bs.FactoryAssociation.greedy(bs.ScoreAssociateEuclideanSq_F64(), float(np.finfo(np.float64).max), True, device=dev.index or 0)
    a.setMode(1)
    p, _ = a.associateDevice(src, dst)                          # warm-up
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        p, _ = a.associateDevice(src, dst)
    e1.record(); torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / reps
    pk, _ = peaks()
    tflops = 2 * 2.0 * n * n * 192 / (ms / 1e3) / 1e12           # two GEMM passes (forward, backward), K = 192
    _, _, (runs, fallbacks) = a.filterScores(src[:1].cpu().numpy(), dst[:1].cpu().numpy(), want_scores=False)
    out = {"workload": "%d x %d SURF-64 descriptors, greedy + backward validation, HBM resident" % (n, n), "ms": ms,
           "pairs_per_s": n * n / (ms / 1e3), "tensor_tflops_whole_call": tflops, "frac_of_bf16_peak": tflops / pk["bf16_tflops"],
           "planted_permutation_recovered": bool(torch.equal(p.long(), perm)), "exact_fallbacks": fallbacks}
    a.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--batch", type=int, default=1024, help="1080p frames per step per GPU (BASELINE configs[1]: 1024)")
    ap.add_argument("--chunk", type=int, default=0, help="frames per kernel wave (0 = library default)")
    ap.add_argument("--ways", type=int, default=3, help="streams a device-resident batch is spread over (1..4)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-assoc", action="store_true", help="skip the association leg (100k x 100k, rank 0)")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "native" else args.warmup
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    saved_stdout = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        # NCCL (or the launcher's environment) prints a version banner on stdout when the first communicator comes up: everything
        # before the result goes to stderr at the file-descriptor level, stdout carries the one JSON line
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    g = run_gpu(args, rank, world, local_rank)
    dev = g["dev"]
    ms = torch.tensor([g["ms"]], device=dev, dtype=torch.float64)
    feats = torch.tensor([float(g["total_features"])], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(feats, op=dist.ReduceOp.SUM)
    ms = float(ms.item())
    B = g["B"]
    value = world * B * args.steps / (ms / 1000.0)

    e2e = None
    if not args.no_e2e:
        fps, h2d, d2h = run_e2e(g, max(1, min(args.steps, 2)), world)
        t = torch.tensor([fps], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MIN)          # slowest rank bounds the job
        e2e = {"value": float(t.item()) * world, "unit": UNIT, "h2d_bytes_per_step": int(h2d) * world,
               "d2h_bytes_per_step": int(d2h) * world, "batch_per_gpu": min(B, max(2048 // world, 64))}

    if rank == 0:
        pk, pk_kind = peaks()
        tm = g["timing"]
        stage_ms = {k: v[0] for k, v in tm.items()}
        total_stage = sum(stage_ms.values()) or 1.0
        hess_ms, hess_calls = tm["fast_hessian"]
        chunk_frames = B * args.steps / max(hess_calls, 1)
        # algorithmic bytes of one Fast-Hessian launch: the integral image of its frames, read once (the
        # intensity layers are avoidable traffic and are not counted)
        alg_bytes = chunk_frames * 4.0 * W * H
        achieved = alg_bytes / (hess_ms / max(hess_calls, 1) / 1000.0) / 1e9 if hess_ms > 0 else 0.0
        nfeat = float(feats.item()) / (world * B)
        pipeline_bytes = 3 * 4.0 * W * H + 545.0 * nfeat
        # DRAM bytes of the dominant kernel from the committed ncu --set full capture (profiles/traffic.json), per launch
        traffic, kernel_name = None, "hessian_tile_kernel<1,9,6,4,64,16>"
        try:
            tj = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "traffic.json")))
            traffic = float(tj["fast_hessian_dram_bytes_per_frame"]) * chunk_frames
            kernel_name = tj.get("kernel", kernel_name)
        except Exception:
            pass
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "surfFast detect+describe, %d x 1920x1080 GrayF32 per GPU per step (BASELINE configs[1])" % B,
                       "batch_per_gpu": B, "features_per_frame": nfeat,
                       "ways": args.ways,
                       "l2": "inputs larger than L2 (%.1f GB batch per step)" % (B * W * H * 4 / 1e9)},
            "roofline": {"bound": "hbm", "kernel": kernel_name, "achieved": achieved, "peak": pk["hbm_gbs"],
                         "unit": "GB/s", "frac": achieved / pk["hbm_gbs"], "traffic": traffic, "peak_kind": pk_kind,
                         "frames_per_launch": chunk_frames, "algorithmic_bytes_per_launch": alg_bytes,
                         "share_of_step": hess_ms / total_stage,
                         "timed": "event pair per stage launch over the same %d steps run on one stream (%.1f ms per step); "
                                  "the timed region of `value` overlaps %d streams" % (args.steps, g["ms_single"] / args.steps, args.ways),
                         "pipeline_achieved_gbs": value / world * pipeline_bytes / 1e9,
                         "pipeline_frac": value / world * pipeline_bytes / 1e9 / pk["hbm_gbs"]},
            "stage_ms_per_step": {k: v / args.steps for k, v in stage_ms.items()},
            "gpu_launches": int(g["launches"]),
            "clocks": g["clocks"],
        }
        if e2e:
            out["e2e"] = e2e
        if not args.no_assoc:
            try:
                out["association"] = run_association(g["torch"], dev)
            except Exception as ex:                              # never lose the headline line over the extra leg
                out["association"] = {"error": str(ex)[:200]}
        if not args.no_cpu and world == 1:
            from boofcv_b200.synth import blob_image
            cores = os.cpu_count() or 1
            frames = [blob_image(W, H, 234 + i) for i in range(2)]
            fps, n, _ = cpu_surf_fps(frames, cores, min_seconds=12.0, max_frames=4096)
            out["cpu_baseline"] = {"value": fps, "unit": UNIT, "cores": cores, "kind": "port",
                                   "sample": "%d x 1080p frames (same generator, seeds 234..235 cycled), surfFast, %d threads; "
                                             "C++ restatement of the reference Java (no JVM on this box)" % (n, cores)}
        if saved_stdout is not None:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
