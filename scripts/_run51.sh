set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_select.py tests/test_gpu_compact.py tests/test_zz_no_alloc.py tests/test_gpu_seeds.py tests/test_gpu_correlate.py -x -q -m gpu > gpurun_out/r2ap_tests.log 2>&1
echo "rc $?" >> gpurun_out/r2ap_tests.log
