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
        fps, n, feats = cpu_surf_fps(frames, cores, min_seconds=4.0, max_frames=16)
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
