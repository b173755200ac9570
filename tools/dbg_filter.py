"""Debug aid: mode-1 association against the exact kernel on the smallest failing shape; prints the rows that differ."""
import numpy as np, sys
sys.path.insert(0, "/root/repo")
import boofcv_b200 as bs
DMAX = float(np.finfo(np.float64).max)
def unit(rng, n):
    x = rng.normal(size=(n, 64)); return x / np.linalg.norm(x, axis=1, keepdims=True)
for (n, m, backwards) in ((300, 517, False), (300, 517, True), (1000, 777, False)):
    rng = np.random.default_rng(n + m)
    dst = unit(rng, m); src = unit(rng, n)
    k = min(n, m) // 2
    src[:k] = dst[rng.permutation(m)[:k]] + 0.02 * rng.normal(size=(k, 64))
    res = []
    for mode in (1, 0):
        g = bs.FactoryAssociation.greedy(bs.ScoreAssociateEuclideanSq_F64(), DMAX, backwards); g.setMode(mode)
        g.setSource(src); g.setDestination(dst); g.associate()
        res.append((g.getPairs().copy(), g.getFitQuality().copy()))
        if mode == 1:
            sc, cnt, st = g.filterScores(src, dst)
        g.close()
    bad = np.nonzero(res[0][0] != res[1][0])[0]
    print((n, m, backwards), "rows that differ:", len(bad), bad[:20], "stats", st, "cnt max", cnt.max())
    ex = ((src[:, None, :] - dst[None, :, :]) ** 2).sum(-1)
    for i in bad[:6]:
        j1, j0 = res[0][0][i], res[1][0][i]
        print("   row", i, "tc ->", j1, "exact ->", j0, "d(tc)", ex[i, j1] if j1 >= 0 else None, "d(exact)", ex[i, j0] if j0 >= 0 else None, "row min at", ex[i].argmin(), ex[i].min(),
              "col tile / in-tile", (j0 // 112, j0 % 112) if j0 >= 0 else None, "est err", abs(sc[i, ex[i].argmin()] - ex[i].min()))
