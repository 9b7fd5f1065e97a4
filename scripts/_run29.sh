set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_knn.py -x -q -m gpu > gpurun_out/r2z_t_knn.log 2>&1
echo "rc $?" >> gpurun_out/r2z_t_knn.log
timeout 600 python scripts/resident_steps.py --steps 12 --times > gpurun_out/r2z_steps.log 2>&1
timeout 900 python -m pytest tests/test_gpu_bench_pile.py -x -q -m gpu -k "graph_matches_oracle or selected_genes" > gpurun_out/r2z_t_bench_pile.log 2>&1
echo "rc $?" >> gpurun_out/r2z_t_bench_pile.log
