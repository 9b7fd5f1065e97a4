set -x
mkdir -p gpurun_out
(nvidia-smi topo -m; free -g; lscpu | head -25; numactl -H 2>/dev/null; nproc) > gpurun_out/r2_box.txt 2>&1
MCB_WRITE_GOLDEN=1 timeout 1500 python -m pytest tests/test_gpu_bench_pile.py tests/test_gpu_log_fraction.py -x -q -s -m gpu > gpurun_out/r2_t1.log 2>&1
echo "rc $?" >> gpurun_out/r2_t1.log
timeout 1200 python -m pytest tests -x -q -m gpu --deselect tests/test_gpu_bench_pile.py --durations=15 > gpurun_out/r2_t_all.log 2>&1
echo "rc $?" >> gpurun_out/r2_t_all.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench_a.json 2> gpurun_out/r2_bench_a.err
echo "rc $?" >> gpurun_out/r2_bench_a.err
timeout 600 python scripts/resident_steps.py --steps 3 > gpurun_out/r2_resident.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:downsample_warp|rank_matrix_bucket|balance_|row_cuts|mirror_prune|log_fraction|prepare_rows' --launch-skip 24 -c 13 -o gpurun_out/r2_ncu_step python scripts/resident_steps.py --steps 3 > gpurun_out/r2_ncu_step.log 2>&1
ls -la gpurun_out
