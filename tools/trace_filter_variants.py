import ctypes as C, sys, os, shutil, numpy as np
sys.path.insert(0, "/root/repo")
variant = sys.argv[1]
import boofcv_b200._native as N
N.SO_PATH = os.path.join(os.path.dirname(N.__file__), "libboofsurf_x%s.so" % variant)
import torch, boofcv_b200 as bs
rng = np.random.default_rng(0)
n = 100000
x = rng.normal(size=(n, 64)); x /= np.linalg.norm(x, axis=1, keepdims=True)
y = x[rng.permutation(n)] + 0.02 * rng.normal(size=(n, 64)); y /= np.linalg.norm(y, axis=1, keepdims=True)
g = bs.FactoryAssociation.greedy(bs.ScoreAssociateEuclideanSq_F64(), float(np.finfo(np.float64).max), True); g.setMode(1)
a, b = torch.from_numpy(y).cuda(), torch.from_numpy(x).cuda()
for _ in range(2):
    try: g.associateDevice(a, b)
    except Exception as e: print("assoc:", str(e)[:80])
torch.cuda.synchronize()
lib = N.lib()
out = np.zeros(32 * 2 * 8, np.int64); out2 = np.zeros(32 * 8, np.int64)
lib.bs_debug_tc_trace.argtypes = [C.c_void_p]; lib.bs_debug_tc_trace2.argtypes = [C.c_void_p]
lib.bs_debug_tc_trace(out.ctypes.data); lib.bs_debug_tc_trace2(out2.ctypes.data)
t = out.reshape(32, 2, 8); r = out2.reshape(32, 8)
print("variant NOLD=%s: relay period %.0f   B wait %.0f   slot h0 wait %.0f   publish %.0f   slot h1 wait %.0f   publish %.0f" % (variant, np.diff(r[:, 0]).mean(), (r[:, 1] - r[:, 0]).mean(), (r[:, 2] - r[:, 1]).mean(), (r[:, 3] - r[:, 2]).mean(), (r[:, 4] - r[:, 3]).mean(), (r[:, 5] - r[:, 4]).mean()))
for h in (0, 1):
    e = t[:, h, :]
    print("   half %d: period %.0f  MMAs+commit %.0f  commit->seen %.0f  read %.0f  published->seen by issuer %.0f" % (h, np.diff(e[:, 0]).mean(), (e[:, 1] - e[:, 0]).mean(), (e[:, 3] - e[:, 1]).mean(), (e[:, 4] - e[:, 3]).mean(), (e[:, 0] - r[:, 3 + 2 * h]).mean()))
