#!/usr/bin/env python
"""bench.py — BASELINE.json's metric: 1080p SURF detect+describe frames/s (surfFast, synthetic frames).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--batch B]

A step = one pass of the hot path (integral -> Fast-Hessian -> NMS/refine -> orientation -> SURF-64) over one batch of
1080p GrayF32 frames (BASELINE.json configs[1]: 1024 frames on one B200).  N>1 (torchrun): every rank runs its own batch on
its own GPU — frames are independent, there is no collective on the data path (weak scaling); ranks only meet at the
timing barrier.

  value      frames/s with the batch already resident in HBM (CUDA events, max over ranks)
  e2e        frames/s with HOST (pinned) frames in and HOST feature lists out, copies inside the timed region
  roofline   the largest stage by CUDA-event time (today orientation + descriptor): algorithmic bytes per launch / its
             event time vs the measured HBM peak, with every other stage's figures beside it (roofline.stages)
  association / config3 / config5   the other BASELINE configs, short legs; under torchrun the 100k x 100k association shards
             its source rows over the ranks (one NCCL all-gather of the column partials)
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
    copies of one sub-batch with the kernels of the others.  Returns (frames/s, h2d, d2h bytes per step, parity probe)."""
    torch, det, dev = g["torch"], g["det"], g["dev"]
    from boofcv_b200 import HostFeatureBuffers
    # host memory is bounded when several ranks share one box: 1024 frames are 8.5 GB of pinned input + 6 GB of pinned output per
    # rank, so the ranks of a box share 4096 frames between them
    B = min(g["B"], max(4096 // world, 64))
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
    # Parity probe (outside the timed region): three frames of what the timed call just returned, against the CPU oracle, bit for bit
    probe = "skipped"
    if g.get("rank", 0) == 0:
        from oracle import bsoracle as orc
        orc.build()
        bad = []
        for f in sorted({0, B // 2, B - 1}):
            o = orc.surf_fast(host[f].numpy(), nthreads=os.cpu_count() or 1)
            b, e = int(out.frame_offset[f]), int(out.frame_offset[f + 1])
            ok = (e - b == len(o["x"]) and np.array_equal(out.xy[b:e, 0], o["x"]) and np.array_equal(out.xy[b:e, 1], o["y"])
                  and np.array_equal(out.scale[b:e], o["scale"]) and np.array_equal(out.angle[b:e], o["angle"])
                  and np.array_equal(out.desc[b:e], o["desc"]) and np.array_equal(out.white[b:e].astype(bool), o["white"]))
            if not ok:
                bad.append(f)
        probe = "ok" if not bad else "MISMATCH in frames %s" % bad
    out.close()
    return B * steps / dt, h2d, d2h, probe, B


def _exact_sq(a, b):
    """DescriptorDistance.euclideanSq of every row of a against every row of b: sequential over k, no FMA (numpy, fp64)."""
    s = np.zeros((len(a), len(b)))
    for k in range(a.shape[1]):
        d = a[:, k][:, None] - b[:, k][None, :]
        s = s + d * d
    return s


def association_probe(src, dst, pairs, fit, rows):
    """Exactness of sampled rows of a 100k x 100k result against the reference's rule, evaluated on the CPU in fp64:
    forward = the LAST column with the smallest distance (AssociateGreedy.java:80-91), kept only if no OTHER source row is as
    close or closer to that column (:98-117); fit = that distance bit for bit."""
    a = src[rows]
    w = _exact_sq(a, dst)                                        # [rows][m]
    best = w.min(axis=1)
    jstar = w.shape[1] - 1 - np.argmin(w[:, ::-1], axis=1)       # last index of the minimum
    col = _exact_sq(dst[jstar], src)                             # [rows][n]: distances of the chosen column to every source row
    bad = 0
    for r, i in enumerate(rows):
        c = col[r]
        alone = (c <= best[r]).sum() == 1                        # only row i itself attains (or beats) it
        want_pair, want_fit = (int(jstar[r]), best[r]) if alone else (-1, np.finfo(np.float64).max)
        if int(pairs[i]) != want_pair or float(fit[i]) != want_fit:
            bad += 1
    return bad


def run_association(torch, dist, dev, rank, world, n=100000, reps=5):
    """Second headline of the path (BASELINE north_star, configs[3]): greedy association of n x n SURF-64 descriptors with backward
    validation, descriptors resident in HBM.  Tensor-core filter + exact fp64 re-score (bs_assoc mode 1).  With several ranks
    the SOURCE ROWS are sharded (boofcv_b200.parallel.ShardedAssociate: one all-gather of the column partials over NCCL)."""
    import boofcv_b200 as bs
    from boofcv_b200 import parallel
    g = torch.Generator(device=dev); g.manual_seed(99)              # the same descriptors on every rank
    dst = torch.randn((n, 64), generator=g, device=dev, dtype=torch.float64); dst /= dst.norm(dim=1, keepdim=True)
    perm = torch.randperm(n, generator=g, device=dev)
    src = dst[perm] + 0.02 * torch.randn((n, 64), generator=g, device=dev, dtype=torch.float64)
    src /= src.norm(dim=1, keepdim=True)                       # SURF descriptors are unit vectors (UtilFeature.normalizeL2)
    a = bs.FactoryAssociation.greedy(bs.ScoreAssociateEuclideanSq_F64(), float(np.finfo(np.float64).max), True, device=dev.index or 0)
    a.setMode(1)
    b0, e0 = parallel.shard_range(n, rank, world)
    mine = src[b0:e0].contiguous()
    sa = parallel.ShardedAssociate(parallel.GpuAssocBackend(a), True)
    call = (lambda: sa.associate(mine, dst, row_offset=b0)) if world > 1 else (lambda: a.associateDevice(src, dst))
    p, f = call()                                               # warm-up
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    e_0, e_1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e_0.record()
    for _ in range(reps):
        p, f = call()
    e_1.record(); torch.cuda.synchronize(dev)
    t = torch.tensor([e_0.elapsed_time(e_1) / reps], device=dev, dtype=torch.float64)
    ok = torch.tensor([1 if torch.equal(p.long(), perm[b0:e0]) else 0], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    ms = float(t.item())
    _, _, (runs, fallbacks) = a.filterScores(src[:1].cpu().numpy(), dst[:1].cpu().numpy(), want_scores=False)
    out = None
    if rank == 0:
        pk, _ = peaks()
        alg = 2.0 * n * n * 64 / (ms / 1e3) / 1e12               # the -2 A B^T term, once
        probe_rows = np.random.default_rng(5).choice(e0 - b0, 48, replace=False)
        bad = association_probe(src.cpu().numpy(), dst.cpu().numpy(), p.cpu().numpy(), f.cpu().numpy(), probe_rows) if world == 1 else None
        out = {"workload": "%d x %d SURF-64 descriptors, greedy + backward validation, HBM resident%s" % (
                   n, n, ", source rows sharded over %d ranks" % world if world > 1 else ""),
               "ms": ms, "pairs_per_s": n * float(n) / (ms / 1e3),
               "algorithmic_tflops": alg, "frac_of_bf16_peak": alg / (world * pk["bf16_tflops"]),             # of the peak of ALL the GPUs used
               "frac_of_bf16_peak_sustained": alg / (world * pk.get("bf16_tflops_sustained", pk["bf16_tflops"])),
               "frac_of_one_gpu_bf16_peak": alg / pk["bf16_tflops"],
               "target_ms_for_half_of_peak": 2.0 * n * n * 64 / (0.5 * world * pk["bf16_tflops"] * 1e12) * 1e3,
               "tensor_pipe_tflops_executed": 2 * alg,            # two K = 64 GEMM passes (forward, backward) are issued
               "planted_permutation_recovered": bool(ok.item() == 1), "exact_fallbacks": fallbacks,
               "parity_probe": None if bad is None else ("ok: 48 sampled rows exact vs fp64 CPU rule" if bad == 0 else "%d of 48 sampled rows differ" % bad)}
    a.close()
    del src, dst
    torch.cuda.empty_cache()
    return out


def run_config3(torch, dist, dev, rank, world, local_rank, frames_per_rank=32):
    """BASELINE configs[2]: surfStable (4 octaves, sliding-window orientation, SurfMod descriptor) on 3840x2160 frames, frames
    sharded over the ranks, no collective (a bounded sample: frames_per_rank per GPU)."""
    import boofcv_b200 as bs
    from boofcv_b200.synth import blob_images_torch
    B = frames_per_rank
    frames = blob_images_torch(3840, 2160, B, seed=234 + rank * B, device=dev)
    det = bs.FactoryDetectDescribe.surfStable(None, None, None, np.float32, device=local_rank)
    det.detectBatch(frames); det.totalFeatures()                # warm-up
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    det.detectBatch(frames)
    feats = det.totalFeatures()
    e1.record(); torch.cuda.synchronize(dev)
    t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    fs = torch.tensor([float(feats)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX); dist.all_reduce(fs)
    det.close(); del frames
    torch.cuda.empty_cache()
    ms = float(t.item())
    return {"workload": "surfStable detect+describe, %d x 3840x2160 GrayF32 per GPU (sample of BASELINE configs[2]), frames sharded" % B,
            "frames_per_s": world * B / (ms / 1e3), "ms": ms, "features_per_frame": float(fs.item()) / (world * B), "scaling": "weak"}


def run_config5(torch, dist, dev, rank, world, local_rank, pairs_total=8192, pairs_cap=256):
    """BASELINE configs[4]: stereo-pair stream — per pair detect+describe both 1080p frames and associate left -> right with backward
    validation, everything on the device; pairs sharded over the ranks (8192 / world per rank, capped to a stated sample)."""
    import boofcv_b200 as bs
    from boofcv_b200.synth import blob_images_torch
    P = min(pairs_total // world, pairs_cap)
    left = blob_images_torch(1920 + 16, 1080, P, seed=5000 + rank * P, device=dev)
    frames = torch.empty((2 * P, 1080, 1920), device=dev, dtype=torch.float32)
    frames[0::2] = left[:, :, :1920]                            # right = left shifted by a disparity of 16 px
    frames[1::2] = left[:, :, 16:]
    del left
    det = bs.FactoryDetectDescribe.surfFast(None, None, None, np.float32, device=local_rank)
    assoc = bs.FactoryAssociation.greedy(bs.ScoreAssociateEuclideanSq_F64(), 0.1, True, device=local_rank)
    sub = 32

    def stream():
        matched = 0
        for p0 in range(0, P, sub):
            det.detectBatch(frames[2 * p0:2 * min(p0 + sub, P)])
            k = min(sub, P - p0)
            pairs, _, off = assoc.associateFramesBatch(det, np.arange(0, 2 * k, 2), np.arange(1, 2 * k, 2))
            if p0 == 0:
                matched = int((pairs[:off[1]] >= 0).sum().item())
        return matched

    stream()                                                    # warm-up
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    matched = stream()
    e1.record(); torch.cuda.synchronize(dev)
    t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    _, _, (tc_runs, tc_fallbacks) = assoc.filterScores(np.zeros((1, 64)), np.zeros((1, 64)), want_scores=False)
    det.close(); assoc.close(); del frames
    torch.cuda.empty_cache()
    ms = float(t.item())
    return {"association_tensor_core_runs": int(tc_runs) - 1, "association_exact_fallbacks": int(tc_fallbacks),
            "workload": "stereo stream: %d x (2 x 1920x1080) per GPU (sample of BASELINE configs[4]: 8192 pairs over the ranks), surfFast x2 + "
                        "greedy(backward, maxFit 0.1) per pair, device resident" % P,
            "pairs_per_s": world * P / (ms / 1e3), "frames_per_s": 2 * world * P / (ms / 1e3), "ms": ms, "matches_first_pair": matched,
            "scaling": "weak"}


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
    ap.add_argument("--no-assoc", action="store_true", help="skip the association leg (100k x 100k, rows sharded over the ranks)")
    ap.add_argument("--no-configs", action="store_true", help="skip the short config-3 / config-5 legs")
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
    g["rank"] = rank
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
        fps, h2d, d2h, probe, eb = run_e2e(g, max(1, args.steps), world)
        t = torch.tensor([fps], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MIN)          # slowest rank bounds the job
        rate = float(t.item())
        e2e = {"value": rate * world, "unit": UNIT, "h2d_bytes_per_step": int(h2d) * world,
               "d2h_bytes_per_step": int(d2h) * world, "batch_per_gpu": eb, "steps": max(1, args.steps),
               "h2d_gbs_per_gpu": rate * h2d / eb / 1e9, "d2h_gbs_per_gpu": rate * d2h / eb / 1e9, "parity_probe": probe}
    # the device batch of the headline is no longer needed: free it before the other legs allocate theirs
    det = g["det"]
    timing, ms_single, clocks, launches = g["timing"], g["ms_single"], g["clocks"], g["launches"]
    det.close()
    g.clear()
    torch.cuda.empty_cache()

    extra = {}
    if not args.no_assoc:
        try:
            extra["association"] = run_association(torch, dist, dev, rank, world)
        except Exception as ex:                                  # never lose the headline line over an extra leg
            extra["association"] = {"error": str(ex)[:300]}
    if not args.no_configs:
        for name, fn in (("config3", run_config3), ("config5", run_config5)):
            try:
                extra[name] = fn(torch, dist, dev, rank, world, local_rank)
            except Exception as ex:
                extra[name] = {"error": str(ex)[:300]}

    if rank == 0:
        pk, pk_kind = peaks()
        tm = timing
        steps = args.steps
        frames_timed = float(B * steps)
        nfeat = float(feats.item()) / (world * B)
        # Algorithmic bytes per frame of each stage: integral = read input + write integral; Fast-Hessian + NMS = the
        # integral read once (the intensity layers are avoidable traffic, NMS adds nothing when fused); descriptors = 545 B per feature
        alg = {"integral": 8.0 * W * H, "fast_hessian": 4.0 * W * H, "nms_refine": 0.0, "compaction": 0.0, "orient_describe": 545.0 * nfeat}
        kernels = {"integral": "integral_rows_kernel + integral_cols_kernel", "fast_hessian": "hessian_tma_kernel<1,..> + hessian_tma_kernel<2,..> + hessian_generic_kernel",
                   "nms_refine": "nms_kernel", "compaction": "scan_counts + frame_offsets + gather_points", "orient_describe": "describe_kernel<float>"}
        traffic_per_frame = {}
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
            traffic_per_frame = tj.get("dram_bytes_per_frame", {})
        except Exception:
            pass
        stage_ms = {k: v[0] for k, v in tm.items()}
        total_stage = sum(stage_ms.values()) or 1.0
        stages = {}
        for k, (msk, calls) in tm.items():
            us_frame = msk * 1000.0 / frames_timed
            gbs = alg[k] / (us_frame * 1e-6) / 1e9 if us_frame > 0 else 0.0
            stages[k] = {"us_per_frame": us_frame, "share_of_step": msk / total_stage, "algorithmic_bytes_per_frame": alg[k],
                         "achieved_gbs": gbs, "frac": gbs / pk["hbm_gbs"], "dram_bytes_per_frame_ncu": traffic_per_frame.get(k)}
        dom = max(stage_ms, key=lambda k: stage_ms[k])
        dom_ms, dom_calls = tm[dom]
        frames_per_launch = frames_timed / max(dom_calls, 1)
        alg_launch = alg[dom] * frames_per_launch
        achieved = alg_launch / (dom_ms / max(dom_calls, 1) / 1000.0) / 1e9 if dom_ms > 0 else 0.0
        pipeline_bytes = 3 * 4.0 * W * H + 545.0 * nfeat
        tr = traffic_per_frame.get(dom)
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "surfFast detect+describe, %d x 1920x1080 GrayF32 per GPU per step (BASELINE configs[1])" % B,
                       "batch_per_gpu": B, "features_per_frame": nfeat,
                       "ways": args.ways,
                       "l2": "inputs larger than L2 (%.1f GB batch per step)" % (B * W * H * 4 / 1e9)},
            "roofline": {"bound": "hbm", "kernel": kernels[dom], "stage": dom, "achieved": achieved, "peak": pk["hbm_gbs"],
                         "unit": "GB/s", "frac": achieved / pk["hbm_gbs"], "traffic": (tr * frames_per_launch) if tr else None,
                         "peak_kind": pk_kind, "frames_per_launch": frames_per_launch, "algorithmic_bytes_per_launch": alg_launch,
                         "share_of_step": dom_ms / total_stage,
                         "note": "the largest stage by CUDA-event time; no stage of this path is HBM-bound (L1 / shared-memory load issue "
                                 "limits them, see profiles/): the fractions say how far each is from the bytes it has to move",
                         "stages": stages,
                         "timed": "event pair per stage launch over the same %d steps run on one stream (%.1f ms per step); "
                                  "the timed region of `value` overlaps %d streams" % (args.steps, ms_single / args.steps, args.ways),
                         "pipeline_achieved_gbs": value / world * pipeline_bytes / 1e9,
                         "pipeline_frac": value / world * pipeline_bytes / 1e9 / pk["hbm_gbs"]},
            "stage_ms_per_step": {k: v / args.steps for k, v in stage_ms.items()},
            "gpu_launches": int(launches),
            "clocks": clocks,
        }
        if e2e:
            out["e2e"] = e2e
            out["parity_probe"] = e2e["parity_probe"]
        out.update(extra)
        if not args.no_cpu and world == 1:
            from boofcv_b200.synth import blob_image
            cores = os.cpu_count() or 1
            frames = [blob_image(W, H, 234 + i) for i in range(2)]
            fps1, n1, _ = cpu_surf_fps(frames, 1, min_seconds=3.0, max_frames=64)
            fps, n, _ = cpu_surf_fps(frames, cores, min_seconds=10.0, max_frames=4096)
            out["cpu_baseline"] = {"value": fps, "unit": UNIT, "cores": cores, "kind": "port", "single_thread_value": fps1,
                                   "sample": "%d x 1080p frames on %d threads (+ %d frames on one thread; same generator, seeds 234..235 cycled), "
                                             "surfFast; C++ restatement of the reference Java (no JVM on this box), persistent thread pool with "
                                             "contiguous blocks like BoofConcurrency's ForkJoinPool; the integral image, the scale-space "
                                             "maximum search and the list handling are serial in the reference too, which bounds the scaling "
                                             "with threads" % (n, cores, n1)}
        if saved_stdout is not None:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
