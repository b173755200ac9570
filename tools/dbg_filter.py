import numpy as np, sys
sys.path.insert(0, "/root/repo")
import boofcv_b200 as bs
rng = np.random.default_rng(1)
def unit(n):
    x = rng.normal(size=(n, 64)); return x / np.linalg.norm(x, axis=1, keepdims=True)
a = unit(300); b = unit(517)
g = bs.FactoryAssociation.greedy(bs.ScoreAssociateEuclideanSq_F64(), float(np.finfo(np.float64).max), False); g.setMode(1)
for rep in range(3):
    scores, cnt, _ = g.filterScores(a, b)
    s = np.zeros((300, 517))
    for k in range(64):
        d = a[:, k][:, None] - b[:, k][None, :]; s += d * d
    err = np.abs(scores - s)
    bad = err > 0.01
    print("rep", rep, "bad", bad.sum(), "rows", np.unique(np.nonzero(bad)[0])[:20], "... n rows", len(np.unique(np.nonzero(bad)[0])))
    cols = np.nonzero(bad)[1]
    print(" col tiles", np.unique(cols // 128), "col halves", np.unique((cols % 128) // 64), "row halves", np.unique((np.nonzero(bad)[0] % 256) // 128))
