"""Host <-> device copy bandwidth with EVERY GPU of the box busy at once (one process per GPU under torchrun): the ceiling of
bench.py's end-to-end (`e2e`) leg at N GPUs.  Prints per-GPU and aggregate GB/s for H2D alone, D2H alone and both at once.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/probe/pcie_bw_all.py"""
import json, os, time
import torch, torch.distributed as dist
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = 1024**3
h = torch.empty(n, dtype=torch.uint8, pin_memory=True); d = torch.empty(n, dtype=torch.uint8, device="cuda")
h2 = torch.empty(n, dtype=torch.uint8, pin_memory=True); d2 = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def timed(fn, reps=4):
    fn(); torch.cuda.synchronize()
    if world > 1: dist.barrier()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / reps
    t = torch.tensor([dt], device="cuda", dtype=torch.float64)
    if world > 1: dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
def both():
    with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
a = timed(lambda: d.copy_(h, non_blocking=True)); b = timed(lambda: h2.copy_(d2, non_blocking=True)); c = timed(both)
if rank == 0:
    print(json.dumps({"gpus": world, "bytes_per_copy": n, "h2d_gbs_per_gpu": n / a / 1e9, "d2h_gbs_per_gpu": n / b / 1e9, "both_each_direction_gbs_per_gpu": n / c / 1e9,
                      "h2d_gbs_aggregate": world * n / a / 1e9, "d2h_gbs_aggregate": world * n / b / 1e9, "both_gbs_aggregate_per_direction": world * n / c / 1e9,
                      "cpus": os.cpu_count(), "affinity": len(os.sched_getaffinity(0))}))
if world > 1:
    dist.barrier(); dist.destroy_process_group()
