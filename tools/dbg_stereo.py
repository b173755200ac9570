"""Dev check: association of real SURF descriptors of stereo pairs — time per pair, tensor-core runs and exact fallbacks."""
import sys, os, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import boofcv_b200 as bs
from boofcv_b200.synth import blob_images_torch
P = 16
left = blob_images_torch(1920 + 16, 1080, P, seed=5000, device="cuda")
frames = torch.empty((2 * P, 1080, 1920), device="cuda", dtype=torch.float32)
frames[0::2] = left[:, :, :1920]; frames[1::2] = left[:, :, 16:]
det = bs.FactoryDetectDescribe.surfFast(None, None, None, np.float32)
det.detectBatch(frames); det.totalFeatures()
for mode in (1, 0):
    assoc = bs.FactoryAssociation.greedy(bs.ScoreAssociateEuclideanSq_F64(), 0.1, True); assoc.setMode(mode)
    assoc.associateFramesBatch(det, np.arange(0, 2 * P, 2), np.arange(1, 2 * P, 2))
    torch.cuda.synchronize(); t = time.perf_counter()
    pairs, fit, off = assoc.associateFramesBatch(det, np.arange(0, 2 * P, 2), np.arange(1, 2 * P, 2))
    torch.cuda.synchronize(); dt = time.perf_counter() - t
    _, cnt, (runs, fb) = assoc.filterScores(np.zeros((1, 64)), np.zeros((1, 64)), want_scores=False)
    print("mode", mode, "ms per pair", 1e3 * dt / P, "runs", runs, "fallbacks", fb, "matched first", int((pairs[:off[1]] >= 0).sum()))
    if mode == 1:
        r0, r1 = det.readFrame(0), det.readFrame(1)
        _, cnt, _ = assoc.filterScores(r0["desc"], r1["desc"], want_scores=False)
        print("candidates per row: mean %.1f max %d, rows over 64: %d of %d" % (cnt.mean(), cnt.max(), (cnt > 64).sum(), len(cnt)))
