set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_partitions.py -x -q -m gpu > gpurun_out/r2n_t_part.log 2>&1
echo "rc $?" >> gpurun_out/r2n_t_part.log
timeout 600 python scripts/time_partitions.py --no-cpu --repeat 1 > gpurun_out/r2n_tp16000.log 2>&1
timeout 900 python scripts/time_partitions.py --cells 10000 --k 750 --no-cpu --repeat 1 > gpurun_out/r2n_tp10000k750.log 2>&1
timeout 600 python scripts/time_partitions.py --cells 4000 --no-cpu --repeat 1 > gpurun_out/r2n_tp4000.log 2>&1
