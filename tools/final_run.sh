# The measurement run behind profiles/r2d_*: probes, both bench arms, one full ncu capture of the association filter, the launch list of
# the library's kernels.  Run on the GPU box from the repository root: bash tools/final_run.sh (writes into gpurun_out/).
set -x
timeout 120 tools/probe/hs2_probe > gpurun_out/r2d_hs2_probe.txt 2>&1
timeout 120 tools/probe/alu_probe > gpurun_out/r2d_alu_probe.txt 2>&1
timeout 900 python bench.py > gpurun_out/r2d_bench_default.json 2> gpurun_out/r2d_bench_default.err
timeout 600 python bench.py --impl reference > gpurun_out/r2d_bench_reference.json 2> gpurun_out/r2d_bench_reference.err
ncu --set full --import-source on --clock-control none -k regex:"filter_kernel" -c 2 -o gpurun_out/r2d_assoc_filter_full python tools/bench_assoc.py --sizes 100000 --exact-max 0 --reps 1 > gpurun_out/r2d_ncu_assoc.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"hessian|describe|nms|compact|integral|filter_kernel|rescore|split_kernel|norm_flags|row_bounds|rowmin|merge|sliding|planar|average_bands|orient|count|scatter|offset" -c 4000 --csv --log-file gpurun_out/r2d_launches.csv python bench.py --steps 1 --warmup 3 --batch 128 --no-cpu --no-configs > gpurun_out/r2d_ncu.log 2>&1
tail -c 600 gpurun_out/r2d_bench_default.json
