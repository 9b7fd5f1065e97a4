set -x
mkdir -p gpurun_out
timeout 300 python scripts/time_partitions.py --cells 4000 --no-cpu --repeat 1 > gpurun_out/r2d_tp4000.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:optimize_partitions_kernel -c 1 -o gpurun_out/r2d_ncu_opt python scripts/time_partitions.py --cells 4000 --no-cpu --repeat 1 > gpurun_out/r2d_ncu_opt.log 2>&1
echo "rc $?" >> gpurun_out/r2d_ncu_opt.log
