"""Integral-image stage alone, device-resident batches: us per 1080p frame and fraction of the measured HBM peak on ALGORITHMIC bytes
(input read once + integral written once), for GrayF32 (serial order, bit-exact), GrayU8 and GrayS32 (parallel scan)."""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import boofcv_b200 as bs
W, H, B = 1920, 1080, 256
pk = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists("MEASURED_PEAKS.json") else 6548.8
g = torch.Generator(device="cuda"); g.manual_seed(1)
base = torch.randint(0, 256, (B, H, W), generator=g, device="cuda", dtype=torch.int32)
for name, frames, dt, fast, inb in (("GrayF32 serial (bit-exact)", base.float(), np.float32, False, 4), ("GrayU8 -> S32 scan", base.to(torch.uint8), np.uint8, False, 1), ("GrayS32 scan", base, np.int32, False, 4)):
    det = bs.FactoryDetectDescribe.surfFast(bs.ConfigFastHessian(1e9, 2, -1, 1, 9, 4, 4), None, None, dt)   # threshold so high that nothing is detected
    det.setWays(1)
    det.detectBatch(frames); det.totalFeatures()
    det.setTiming(True)
    for _ in range(3): det.detectBatch(frames)
    det.totalFeatures()
    ms, calls = det.getTiming()["integral"]
    us = ms * 1e3 / (3 * B)
    alg = (inb + 4.0) * W * H
    print(json.dumps({"case": name, "us_per_1080p_frame": us, "algorithmic_bytes_per_frame": alg, "achieved_gbs": alg / us / 1e3, "frac_of_hbm_peak": alg / us / 1e3 / pk}))
    det.close()
