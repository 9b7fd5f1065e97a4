set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_knn.py tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r2ab_t.log 2>&1
echo "rc $?" >> gpurun_out/r2ab_t.log
timeout 600 python scripts/resident_steps.py --steps 12 --times > gpurun_out/r2ab_steps.log 2>&1
timeout 900 python -m pytest tests/test_gpu_bench_pile.py -x -q -m gpu -k "graph_matches_oracle" > gpurun_out/r2ab_t_bench_pile.log 2>&1
echo "rc $?" >> gpurun_out/r2ab_t_bench_pile.log
