"""Dev helper: bench.py at several chunk sizes, one summary line each (run on the GPU box)."""
import json
import subprocess
import sys

batch = sys.argv[1] if len(sys.argv) > 1 else "256"
for c in sys.argv[2:] or ["1", "2", "4", "8"]:
    out = subprocess.run([sys.executable, "bench.py", "--batch", batch, "--steps", "2", "--chunk", c, "--no-cpu", "--no-e2e", "--no-assoc"],
                         capture_output=True, text=True)
    try:
        d = json.loads(out.stdout.strip().splitlines()[-1])
        print("chunk", c, "fps", round(d["value"]), {k: round(v, 2) for k, v in d["stage_ms_per_step"].items()}, flush=True)
    except Exception as e:
        print("chunk", c, "failed", e, out.stderr[-500:], flush=True)
