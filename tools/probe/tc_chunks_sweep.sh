for c in 0 1 2 3 4 6 9; do
BS_TC_CHUNKS=$c ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"filter_kernel" --launch-skip 6 -c 4 --csv --log-file gpurun_out/tcc_$c.csv python tools/prof_assoc_small.py 10000 > /dev/null 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/tcc_$c.csv")) if len(r)>10]
h=rows[0]; vi=h.index("Metric Value"); gi=h.index("Grid Size")
v=[float(r[vi].replace(",","")) for r in rows[1:]]
print("chunks $c grid", rows[1][gi], "filter us", [round(x/1000,1) if x>1000 else x for x in v])
PY
done
