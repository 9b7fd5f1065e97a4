set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_partitions.py -x -q -m gpu > gpurun_out/r2f_t_part.log 2>&1
echo "rc $?" >> gpurun_out/r2f_t_part.log
timeout 600 python scripts/time_partitions.py --no-cpu --repeat 1 > gpurun_out/r2f_tp16000.log 2>&1
timeout 300 python scripts/time_partitions.py --cells 4000 --no-cpu --repeat 1 > gpurun_out/r2f_tp4000.log 2>&1
