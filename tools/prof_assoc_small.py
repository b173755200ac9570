"""Small-problem association timing (10k x 10k, the per-pair size of BASELINE config 5)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, boofcv_b200 as bs
dev = torch.device("cuda")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
g = torch.Generator(device=dev); g.manual_seed(1)
dst = torch.randn((n, 64), generator=g, device=dev, dtype=torch.float64); dst /= dst.norm(dim=1, keepdim=True)
src = dst[torch.randperm(n, generator=g, device=dev)] + 0.02 * torch.randn((n, 64), generator=g, device=dev, dtype=torch.float64)
a = bs.FactoryAssociation.greedy(bs.ScoreAssociateEuclideanSq_F64(), 0.1, True)
for _ in range(3):
    a.associateDevice(src, dst)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(20):
    a.associateDevice(src, dst)
torch.cuda.synchronize()
print("n", n, "ms per call (wall)", (time.perf_counter() - t0) / 20 * 1e3)
