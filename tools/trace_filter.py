"""Timing experiment: clock64 stamps of the accumulator hand-off inside filter_kernel (library built with -DBS_TC_TRACE).
Prints, for 32 consecutive tiles of CTA 0 in steady state and both A halves, the intervals between: first MMA issued (0), commit issued (1),
issuer past its wait for the other slot (2), epilogue group sees acc_full (3), its TMEM read complete (4), slot released (5)."""
import ctypes as C, sys, numpy as np
sys.path.insert(0, "/root/repo")
import torch
import boofcv_b200 as bs
from boofcv_b200 import _native as N
rng = np.random.default_rng(0)
n = 100000
x = rng.normal(size=(n, 64)); x /= np.linalg.norm(x, axis=1, keepdims=True)
y = x[rng.permutation(n)] + 0.02 * rng.normal(size=(n, 64)); y /= np.linalg.norm(y, axis=1, keepdims=True)
g = bs.FactoryAssociation.greedy(bs.ScoreAssociateEuclideanSq_F64(), float(np.finfo(np.float64).max), True); g.setMode(1)
a, b = torch.from_numpy(y).cuda(), torch.from_numpy(x).cuda()
for _ in range(2): g.associateDevice(a, b)
torch.cuda.synchronize()
out = np.zeros(32 * 2 * 8, np.int64)
lib = N.lib()
lib.bs_debug_tc_trace.argtypes = [C.c_void_p]
assert lib.bs_debug_tc_trace(out.ctypes.data) == 0
t = out.reshape(32, 2, 8)          # the LAST launch = backward pass
for h in (0, 1):
    e = t[:, h, :]
    period = np.diff(e[:, 0])
    print("half", h, "tile period (first MMA to first MMA): mean %.0f, min %d, max %d cycles" % (period.mean(), period.min(), period.max()))
    print("   issue of 4 MMAs + commit (0->1): %.0f   commit -> epilogue sees acc_full (1->3): %.0f   TMEM read (3->4): %.0f   release (4->5): %.0f" % (
        (e[:, 1] - e[:, 0]).mean(), (e[:, 3] - e[:, 1]).mean(), (e[:, 4] - e[:, 3]).mean(), (e[:, 5] - e[:, 4]).mean()))
    # slot round trip: tile n's release (5) -> first MMA of tile n + 2 (same slot)
    print("   release of tile n -> first MMA of tile n+2 (same slot): %.0f   issuer wait after commit (1->2): %.0f   probes that succeeded: %d of 32" % (
        (e[2:, 0] - e[:-2, 5]).mean(), (e[:, 2] - e[:, 1]).mean(), int(e[:, 6].sum())))
    print("   slot cycle (first MMA n -> first MMA n+2): %.0f" % (e[2:, 0] - e[:-2, 0]).mean())
