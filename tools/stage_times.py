"""Per-stage device time (us per frame) of surfFast / surfStable at 1080p and 2160p."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, boofcv_b200 as bs
from boofcv_b200.synth import blob_images_torch
which = sys.argv[1:] or ["fast1080", "stable1080", "stable2160"]
cfg = {"fast1080": (bs.FactoryDetectDescribe.surfFast, (1920, 1080), 64), "stable1080": (bs.FactoryDetectDescribe.surfStable, (1920, 1080), 64),
       "stable2160": (bs.FactoryDetectDescribe.surfStable, (3840, 2160), 32), "fast2160": (bs.FactoryDetectDescribe.surfFast, (3840, 2160), 32)}
for name in which:
    fac, (w, h), B = cfg[name]
    fr = blob_images_torch(w, h, B, seed=1, device=torch.device("cuda"))
    det = fac(None, None, None, np.float32)
    det.detectBatch(fr); det.totalFeatures(); det.setTiming(True)
    det.detectBatch(fr); n = det.totalFeatures()
    t = {k: round(v[0] / B * 1000, 1) for k, v in det.getTiming().items()}
    print(name, "features/frame %.0f" % (n / B), t, "total %.1f us/frame" % sum(t.values()))
    det.close(); del fr
