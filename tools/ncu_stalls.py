"""Stall / instruction-mix digest of one kernel out of an ncu --set full --import-source on report.
usage: python tools/ncu_stalls.py report.ncu-rep [top_n] [launch_index]   (launch_index: which launch of a multi-kernel report)"""
import csv
import subprocess
import sys
from collections import Counter

out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"] + [len(rows)]
k = int(sys.argv[3]) if len(sys.argv) > 3 else 0
rows = rows[starts[k]:starts[k + 1]]
print("launch", k, "of", len(starts) - 1, ":", rows[0][1][:100])
h = rows[1]; c = {n: i for i, n in enumerate(h)}
data = rows[2:]
I = lambda r, n: int(r[c[n]] or 0)
cols = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
tot = {n: sum(I(r, n) for r in data) for n in cols}
print("samples", sum(I(r, "# Samples") for r in data), {k: v for k, v in sorted(tot.items(), key=lambda x: -x[1])[:10]})
ldtm = [I(r, "Instructions Executed") for r in data if "LDTM" in r[c["Source"]]]
per = max(ldtm) if ldtm else max(I(r, "Instructions Executed") for r in data)
cnt = Counter()
for r in data:
    op = [o for o in r[c["Source"]].split() if not o.startswith("@")]
    cnt[op[0].split(".")[0] if op else ""] += I(r, "Instructions Executed")
print("instructions per", per, "executions of the hottest LDTM (or line):", [(k, round(v / per, 1)) for k, v in cnt.most_common(24)])
for r in sorted(data, key=lambda r: -I(r, "# Samples"))[:int(sys.argv[2]) if len(sys.argv) > 2 else 25]:
    st = {n[6:]: I(r, n) for n in cols if I(r, n) > 0.15 * max(I(r, "# Samples"), 1)}
    print(r[c["Address"]][-5:], str(I(r, "# Samples")).rjust(6), str(I(r, "Instructions Executed")).rjust(9), r[c["Source"]][:84], st)
