"""Dev helper: can the smem-bound Fast-Hessian kernels and the L1-bound describe kernel share SMs?  Both are persistent and by
default each fills every SM with its own CTAs, so kernels of different streams ("ways") queue behind each other.  This runs
bench.py's device-resident leg under several per-SM CTA caps (BS_HT_PER_SM, BS_DESC_CTAS) and stream counts."""
import itertools, json, os, subprocess, sys
combos = [(4, 7, 3)] + [(ht, dc, w) for ht, dc, w in itertools.product((2, 3), (3, 4, 5), (3, 4))] + [(1, 3, 4), (2, 2, 4)]
for ht, dc, w in combos:
    env = dict(os.environ); env["BS_HT_PER_SM"] = str(ht); env["BS_DESC_CTAS"] = str(dc)
    out = subprocess.run([sys.executable, "bench.py", "--batch", "512", "--steps", "2", "--ways", str(w), "--no-cpu", "--no-e2e", "--no-assoc", "--no-configs"],
                         capture_output=True, text=True, env=env)
    try:
        d = json.loads(out.stdout.strip().splitlines()[-1])
        print("hessian CTAs/SM", ht, "describe CTAs/SM", dc, "ways", w, "fps", round(d["value"]), {k: round(x, 1) for k, x in d["stage_ms_per_step"].items()}, flush=True)
    except Exception as e:
        print(ht, dc, w, "failed", e, out.stderr[-300:], flush=True)
