"""Dev helper: bench.py under several values of one environment variable (run on the GPU box).
usage: sweep_env.py VAR v1 v2 ... [-- extra bench args]"""
import json
import os
import subprocess
import sys

args = sys.argv[1:]
extra = []
if "--" in args:
    k = args.index("--")
    args, extra = args[:k], args[k + 1:]
var, vals = args[0], args[1:]
for v in vals:
    env = dict(os.environ)
    env[var] = v
    out = subprocess.run([sys.executable, "bench.py", "--batch", "256", "--steps", "2", "--chunk", "8", "--no-cpu", "--no-e2e", "--no-assoc"] + extra,
                         capture_output=True, text=True, env=env)
    try:
        d = json.loads(out.stdout.strip().splitlines()[-1])
        print(var, v, "fps", round(d["value"]), {k: round(x, 2) for k, x in d["stage_ms_per_step"].items()}, flush=True)
    except Exception as e:
        print(var, v, "failed", e, out.stderr[-500:], flush=True)
