"""Per-kernel durations out of an `ncu --metrics gpu__time_duration.sum --csv` log: python tools/kernel_times.py log.csv [filter]"""
import csv
import sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
h = rows[0]; ki = h.index("Kernel Name"); vi = h.index("Metric Value")
for r in rows[1:]:
    if len(sys.argv) < 3 or sys.argv[2] in r[ki]:
        print("%-70s %10.1f us" % (r[ki][:70], float(r[vi].replace(",", "")) / 1e3))
