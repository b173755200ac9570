"""Multi-GPU check, run under torchrun on the GPU box (one rank per GPU, NCCL):
  1. row-sharded association (parallel.ShardedAssociate over NCCL, tensor-core filter per rank) == single-GPU result;
  2. frame-sharded detect+describe: every rank's frames equal the same frames processed alone (no collective involved).
Prints one JSON line from rank 0.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/mgpu_check.py"""
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


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    out = {"world": world}
    # ---- association: same data on every rank (seeded), rows split across ranks
    n = 40000
    g = torch.Generator(device=dev); g.manual_seed(7)
    dst = torch.randn((n, 64), generator=g, device=dev, dtype=torch.float64); dst /= dst.norm(dim=1, keepdim=True)
    perm = torch.randperm(n, generator=g, device=dev)
    src = dst[perm] + 0.02 * torch.randn((n, 64), generator=g, device=dev, dtype=torch.float64)
    src[11] = src[10]                                           # a column tie across what may be two shards
    fmax = float(np.finfo(np.float64).max)
    a = bs.FactoryAssociation.greedy(bs.ScoreAssociateEuclideanSq_F64(), fmax, True, device=local)
    a.setMode(1)
    b, e = parallel.shard_range(n, rank, world)
    sa = parallel.ShardedAssociate(parallel.GpuAssocBackend(a), True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    pairs, fit = sa.associate(src[b:e].contiguous(), dst)       # warm-up
    dist.barrier(); torch.cuda.synchronize()
    e0.record()
    pairs, fit = sa.associate(src[b:e].contiguous(), dst)
    e1.record(); torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev); dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    gp, gf = parallel.gather_matches(pairs, fit)
    if rank == 0:
        p1, f1 = a.associateDevice(src, dst)                    # single GPU, all rows
        torch.cuda.synchronize()
        out["assoc_sharded_equals_single_gpu"] = bool(np.array_equal(gp, p1.cpu().numpy()) and np.array_equal(gf, f1.cpu().numpy()))
        out["assoc_matches"] = int((gp >= 0).sum())
        out["assoc_sharded_ms"] = float(ms.item())
        out["assoc_pairs_per_s"] = n * n / (float(ms.item()) / 1e3)
    # ---- frames: rank r takes its slice of a 16-frame batch
    B = 16
    frames = blob_images_torch(640, 480, B, seed=900, device=dev)
    fb, fe = parallel.frame_shard(B, rank, world)
    det = bs.FactoryDetectDescribe.surfFast(None, None, None, np.float32, device=local)
    det.detectBatch(frames[fb:fe])
    mine = [det.readFrame(f) for f in range(fe - fb)]
    ok = True
    for i, f in enumerate(range(fb, fe)):
        det.detectBatch(frames[f:f + 1])
        r = det.readFrame(0)
        ok = ok and all(np.array_equal(r[k], mine[i][k]) for k in ("xy", "scale", "angle", "desc", "white"))
    t = torch.tensor([1 if ok else 0], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MIN)
    cnt = torch.tensor([sum(len(m["x"]) for m in mine)], device=dev); dist.all_reduce(cnt)
    if rank == 0:
        out["frames_sharded_equal_single"] = bool(t.item() == 1)
        out["frames_total_features"] = int(cnt.item())
        print(json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
