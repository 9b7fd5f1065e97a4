set -x
mkdir -p gpurun_out
for W in 2 4; do
  MCB_OPT_TEAM=$W timeout 900 python -m pytest tests/test_gpu_partitions.py -x -q -m gpu > gpurun_out/r2m_t_part_w$W.log 2>&1
  echo "rc $?" >> gpurun_out/r2m_t_part_w$W.log
done
for W in 1 2 4; do
  MCB_OPT_TEAM=$W timeout 600 python scripts/time_partitions.py --no-cpu --repeat 1 > gpurun_out/r2m_tp16000_w$W.log 2>&1
  MCB_OPT_TEAM=$W timeout 900 python scripts/time_partitions.py --cells 10000 --k 750 --no-cpu --repeat 1 > gpurun_out/r2m_tp10000k750_w$W.log 2>&1
done
