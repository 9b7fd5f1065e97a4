set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_limits.py -x -q -m gpu > gpurun_out/r2x_t_limits.log 2>&1
echo "rc $?" >> gpurun_out/r2x_t_limits.log
