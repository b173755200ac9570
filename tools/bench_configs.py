"""BASELINE.json configs 3, 4 and 5 (the multi-GPU cases; bench.py measures configs[1]).  One process per GPU:

  python tools/bench_configs.py                                  # 1 GPU
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29513 tools/bench_configs.py

  config 3  surfStable (4 octaves, sub-pixel refine) on 3840x2160 frames, frames sharded over the ranks (no collective)
  config 4  AssociateGreedy + backward validation, 100k x 100k SURF-64, source rows sharded, NCCL all-gather of the column minima
  config 5  stereo stream: 1080p frame pairs, detect+describe both frames, associate left->right on the device (pairs sharded)

Frame / pair counts are per rank (weak scaling) and smaller than BASELINE's full streams by default (--frames3, --pairs5); rates
are device-timed (CUDA events, max over ranks) after a warm-up pass.  Prints one JSON line per config from rank 0."""
import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import boofcv_b200 as bs  # noqa: E402
from boofcv_b200 import parallel  # noqa: E402
from boofcv_b200.synth import blob_images_torch  # noqa: E402


def timed(fn, dev, world, reps=1):
    fn()                                                           # warm-up (workspaces, first-use costs)
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record(); torch.cuda.synchronize(dev)
    ms = torch.tensor([e0.elapsed_time(e1) / reps], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item()), out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames3", type=int, default=64, help="2160p frames per rank (config 3)")
    ap.add_argument("--n4", type=int, default=100000, help="descriptors per side (config 4)")
    ap.add_argument("--pairs5", type=int, default=128, help="1080p stereo pairs per rank (config 5)")
    ap.add_argument("--only", type=int, default=0)
    ap.add_argument("--pipeline5", action="store_true", help="config 5: two detector / associator handles on two streams (measured slower: "
                    "the persistent kernels fill the SMs, 1530 vs 1705 pairs/s)")
    ap.add_argument("--per-pair-calls", action="store_true", help="config 5: one associate call (and host round trip) per pair")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    say = (lambda d: print(json.dumps(d), flush=True)) if rank == 0 else (lambda d: None)

    if args.only in (0, 3):
        B = args.frames3
        frames = blob_images_torch(3840, 2160, B, seed=234 + rank * B, device=dev)
        det = bs.FactoryDetectDescribe.surfStable(None, None, None, np.float32, device=local)
        ms, _ = timed(lambda: (det.detectBatch(frames), det.totalFeatures())[1], dev, world)
        feats = torch.tensor([float(det.totalFeatures())], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(feats)
        say({"config": 3, "workload": "surfStable detect+describe, %d x 3840x2160 GrayF32 per GPU, frames sharded" % B, "n_gpus": world,
             "frames_per_s": world * B / (ms / 1e3), "ms": ms, "features_per_frame": float(feats.item()) / (world * B), "scaling": "weak"})
        det.close(); del frames, det
        torch.cuda.empty_cache()

    if args.only in (0, 4):
        n = args.n4
        g = torch.Generator(device=dev); g.manual_seed(99)          # same descriptors on every rank
        dst = torch.randn((n, 64), generator=g, device=dev, dtype=torch.float64); dst /= dst.norm(dim=1, keepdim=True)
        perm = torch.randperm(n, generator=g, device=dev)
        src = dst[perm] + 0.02 * torch.randn((n, 64), generator=g, device=dev, dtype=torch.float64)
        a = bs.FactoryAssociation.greedy(bs.ScoreAssociateEuclideanSq_F64(), float(np.finfo(np.float64).max), True, device=local)
        a.setMode(1)
        b, e = parallel.shard_range(n, rank, world)
        mine = src[b:e].contiguous()
        if world > 1:
            sa = parallel.ShardedAssociate(parallel.GpuAssocBackend(a), True)
            ms, (pairs, fit) = timed(lambda: sa.associate(mine, dst), dev, world, reps=3)
        else:
            ms, (pairs, fit) = timed(lambda: a.associateDevice(src, dst), dev, world, reps=3)
        ok = torch.tensor([1 if torch.equal(pairs.long(), perm[b:e]) else 0], device=dev)
        if world > 1:
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        say({"config": 4, "workload": "greedy + backward validation, %d x %d SURF-64, source rows sharded" % (n, n), "n_gpus": world,
             "ms": ms, "pairs_per_s": n * float(n) / (ms / 1e3), "tensor_tflops": 2 * 2.0 * n * n * 192 / (ms / 1e3) / 1e12,
             "planted_permutation_recovered": bool(ok.item() == 1), "scaling": "strong"})
        a.close(); del src, dst
        torch.cuda.empty_cache()

    if args.only in (0, 5):
        P = args.pairs5
        left = blob_images_torch(1920 + 16, 1080, P, seed=5000 + rank * P, device=dev)
        frames = torch.empty((2 * P, 1080, 1920), device=dev, dtype=torch.float32)
        frames[0::2] = left[:, :, :1920]                            # right = left shifted by a disparity of 16 px
        frames[1::2] = left[:, :, 16:]
        del left
        # two detector / associator handles on two streams: the association of sub-batch i (tensor cores) overlaps the detection
        # of sub-batch i+1 (shared-memory and L1 bound); the reference's interfaces allow it (independent instances)
        dets = [bs.FactoryDetectDescribe.surfFast(None, None, None, np.float32, device=local) for _ in range(2)]
        assocs = [bs.FactoryAssociation.greedy(bs.ScoreAssociateEuclideanSq_F64(), 0.1, True, device=local) for _ in range(2)]
        lanes = [torch.cuda.Stream(device=dev) for _ in range(2)]
        det, assoc = dets[0], assocs[0]
        sub = 32                                                    # pairs per detect call: association of one sub-batch follows it

        def stream_pipelined():
            matched = 0
            cur = torch.cuda.current_stream(dev)
            for l in lanes:
                l.wait_stream(cur)
            starts = list(range(0, P, sub))

            def detect(i):
                p0 = starts[i]
                dets[i % 2].detectBatch(frames[2 * p0:2 * min(p0 + sub, P)], stream=lanes[i % 2].cuda_stream)

            detect(0)
            for i, p0 in enumerate(starts):
                if i + 1 < len(starts):
                    detect(i + 1)                                   # queued before the host waits for sub-batch i
                k = min(sub, P - p0)
                pairs, _, off = assocs[i % 2].associateFramesBatch(dets[i % 2], np.arange(0, 2 * k, 2), np.arange(1, 2 * k, 2),
                                                                   stream=lanes[i % 2].cuda_stream)
                if p0 == 0:
                    matched = int((pairs[:off[1]] >= 0).sum().item())
            for l in lanes:
                cur.wait_stream(l)
            return matched

        def stream_serial():
            matched = 0
            for p0 in range(0, P, sub):
                det.detectBatch(frames[2 * p0:2 * min(p0 + sub, P)])
                k = min(sub, P - p0)
                if args.per_pair_calls:
                    for p in range(k):
                        pairs, _ = assoc.associateFrames(det, 2 * p, 2 * p + 1)
                        if p == 0 and p0 == 0:
                            matched = int((pairs >= 0).sum().item())
                else:                                               # one call per sub-batch: pairs queued back to back on the device
                    pairs, _, off = assoc.associateFramesBatch(det, np.arange(0, 2 * k, 2), np.arange(1, 2 * k, 2))
                    if p0 == 0:
                        matched = int((pairs[:off[1]] >= 0).sum().item())
            return matched

        stream = stream_pipelined if (args.pipeline5 and not args.per_pair_calls) else stream_serial
        ms, matched = timed(stream, dev, world)
        say({"config": 5, "workload": "stereo stream: %d x (2 x 1920x1080) per GPU, surfFast x2 + greedy(backward) per pair, device resident" % P,
             "n_gpus": world, "pairs_per_s": world * P / (ms / 1e3), "frames_per_s": 2 * world * P / (ms / 1e3), "ms": ms,
             "matches_first_pair": matched, "scaling": "weak"})
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
