set -x
mkdir -p gpurun_out
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --atlas-cells 0 --candidate-piles 0 > gpurun_out/r02_bench_for_ncu.json 2> gpurun_out/r02_bench_for_ncu.err && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02_ncu_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --atlas-cells 0 --candidate-piles 0 > gpurun_out/r02_ncu_launches.log 2>&1
timeout 300 python scripts/resident_steps.py --steps 3 > gpurun_out/r02_resident.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k 'regex:downsample_warp|rank_matrix_bucket|balance_kernel|row_cuts|mirror_prune|log_fraction|prepare_rows|correlate_umma|sliding_median|gene_columns' --launch-skip 26 -c 14 -o gpurun_out/r02_ncu_step python scripts/resident_steps.py --steps 3 > gpurun_out/r02_ncu_step.log 2>&1
timeout 1200 compute-sanitizer --tool memcheck --print-limit 20 python __graft_entry__.py smoke > gpurun_out/r02_sanitizer_memcheck.log 2>&1
echo "rc $?" >> gpurun_out/r02_sanitizer_memcheck.log
timeout 1200 compute-sanitizer --tool racecheck --print-limit 20 python __graft_entry__.py smoke > gpurun_out/r02_sanitizer_racecheck.log 2>&1
echo "rc $?" >> gpurun_out/r02_sanitizer_racecheck.log
ls -la gpurun_out
