set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_select.py -x -q -m gpu > gpurun_out/r2w_t_select.log 2>&1
echo "rc $?" >> gpurun_out/r2w_t_select.log
timeout 600 python scripts/resident_steps.py --steps 12 --times > gpurun_out/r2w_steps.log 2>&1
timeout 1200 python -m pytest tests/test_gpu_bench_pile.py -x -q -m gpu -k "selected_genes or to_metacells" > gpurun_out/r2w_t_bench_pile.log 2>&1
echo "rc $?" >> gpurun_out/r2w_t_bench_pile.log
