set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_compact.py -x -q -m gpu > gpurun_out/r2y_t_compact.log 2>&1
echo "rc $?" >> gpurun_out/r2y_t_compact.log
timeout 1500 python bench.py --atlas-cells 0 --no-cpu-baseline --candidate-piles 0 > gpurun_out/r2y_bench.json 2> gpurun_out/r2y_bench.err
echo "rc $?" >> gpurun_out/r2y_bench.err
