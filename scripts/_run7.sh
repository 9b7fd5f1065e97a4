set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_select.py tests/test_gpu_downsample.py tests/test_gpu_atlas.py -x -q -m gpu > gpurun_out/r2g_t_select.log 2>&1
echo "rc $?" >> gpurun_out/r2g_t_select.log
timeout 300 python scripts/time_select.py > gpurun_out/r2g_time_select.log 2>&1
echo "rc $?" >> gpurun_out/r2g_time_select.log
