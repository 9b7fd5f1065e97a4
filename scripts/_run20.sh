set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_deviants.py -x -q -m gpu > gpurun_out/r2s_t_dev.log 2>&1
echo "rc $?" >> gpurun_out/r2s_t_dev.log
