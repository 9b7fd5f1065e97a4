set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_select.py -x -q -m gpu > gpurun_out/r2h_t_select.log 2>&1
echo "rc $?" >> gpurun_out/r2h_t_select.log
MCB_WRITE_GOLDEN=1 timeout 900 python -m pytest tests/test_gpu_bench_pile.py -x -q -s -m gpu -k "selected_genes" > gpurun_out/r2h_t_bench_select.log 2>&1
echo "rc $?" >> gpurun_out/r2h_t_bench_select.log
timeout 900 python bench.py --atlas-cells 0 --candidate-piles 0 > gpurun_out/r2h_bench.json 2> gpurun_out/r2h_bench.err
echo "rc $?" >> gpurun_out/r2h_bench.err
