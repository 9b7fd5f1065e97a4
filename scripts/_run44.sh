set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_partitions.py tests/test_gpu_limits.py tests/test_gpu_candidates.py -x -q -m gpu > gpurun_out/r2ai_tests.log 2>&1
echo "rc $?" >> gpurun_out/r2ai_tests.log
timeout 600 python scripts/time_partitions.py --no-cpu > gpurun_out/r2ai_time.log 2>&1
echo "rc $?" >> gpurun_out/r2ai_time.log
