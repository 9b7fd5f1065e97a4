set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_knn.py tests/test_gpu_deviants.py tests/test_gpu_atlas.py -x -q -m gpu > gpurun_out/r2aa_t.log 2>&1
echo "rc $?" >> gpurun_out/r2aa_t.log
timeout 600 python scripts/resident_steps.py --steps 12 --times > gpurun_out/r2aa_steps.log 2>&1
timeout 900 python -m pytest tests/test_gpu_bench_pile.py -x -q -m gpu -k "graph_matches_oracle or selected_genes" > gpurun_out/r2aa_t_bench_pile.log 2>&1
echo "rc $?" >> gpurun_out/r2aa_t_bench_pile.log
timeout 600 python scripts/time_candidates.py --out gpurun_out/r2aa_candidates.json > gpurun_out/r2aa_tc.log 2>&1
