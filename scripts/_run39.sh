set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_atlas.py tests/test_gpu_downsample.py -x -q -m gpu > gpurun_out/r2ae_t_atlas.log 2>&1
echo "rc $?" >> gpurun_out/r2ae_t_atlas.log
