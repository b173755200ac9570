"""Association benchmark (BASELINE config: N x M SURF-64 descriptors, greedy + backward validation), run on the GPU box.

Prints one JSON line per size: pairs/s for the tensor-core filter path (mode 1) and, where it finishes quickly, the exact
fp64 kernel (mode 0); the tensor throughput the filter's GEMM passes reach against the measured bf16 peak; whether both
modes returned identical pairs/fit.  Descriptors are synthetic unit vectors with planted noisy matches, resident in HBM.

usage: bench_assoc.py [--sizes 10000,50000,100000] [--exact-max 50000] [--reps 3]"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="10000,50000,100000")
    ap.add_argument("--exact-max", type=int, default=50000)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--no-backwards", action="store_true")
    args = ap.parse_args()
    import torch
    import boofcv_b200 as bs
    dev = torch.device("cuda:0")
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
    except Exception:
        pass
    bf16_peak = float(peaks.get("bf16_tflops_burst", peaks.get("bf16_tflops", 1656.0)))
    backwards = not args.no_backwards
    for n in [int(s) for s in args.sizes.split(",")]:
        g = torch.Generator(device=dev); g.manual_seed(n)
        dst = torch.randn((n, 64), generator=g, device=dev, dtype=torch.float64)
        dst /= dst.norm(dim=1, keepdim=True)
        perm = torch.randperm(n, generator=g, device=dev)
        src = dst[perm] + 0.02 * torch.randn((n, 64), generator=g, device=dev, dtype=torch.float64)
        src /= src.norm(dim=1, keepdim=True)
        out = {"n": n, "m": n, "dof": 64, "backwards_validation": backwards}
        res = {}
        for mode in (1, 0):
            if mode == 0 and n > args.exact_max:
                continue
            a = bs.FactoryAssociation.greedy(bs.ScoreAssociateEuclideanSq_F64(), float(np.finfo(np.float64).max), backwards)
            a.setMode(mode)
            a.associateDevice(src, dst)                          # warm-up (workspace, tensor maps)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(args.reps):
                p, f = a.associateDevice(src, dst)
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / args.reps
            res[mode] = (p.clone(), f.clone())
            key = "tensor_filter" if mode == 1 else "exact_fp64"
            out[key] = {"ms": ms, "pairs_per_s": n * n / (ms / 1e3)}
            if mode == 1:
                passes = 2 if backwards else 1
                flops = passes * 2.0 * n * n * 192
                out[key]["gemm_tflops_incl_rescore_time"] = flops / (ms / 1e3) / 1e12
                out[key]["frac_of_bf16_peak"] = out[key]["gemm_tflops_incl_rescore_time"] / bf16_peak
                _, _, (runs, fb) = a.filterScores(src[:1].cpu().numpy(), dst[:1].cpu().numpy(), want_scores=False)
                out[key]["fallbacks"] = fb
                out[key]["recovered_planted_permutation"] = bool(torch.equal(p.long(), perm))
            a.close()
        if 0 in res and 1 in res:
            out["modes_identical"] = bool(torch.equal(res[0][0], res[1][0]) and torch.equal(res[0][1], res[1][1]))
            out["speedup_vs_exact"] = out["exact_fp64"]["ms"] / out["tensor_filter"]["ms"]
        out["bf16_peak_tflops"] = bf16_peak
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
