set -x
mkdir -p gpurun_out
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --atlas-cells 0 --candidate-piles 0 > gpurun_out/r02c_bench_for_ncu.json 2> gpurun_out/r02c_bench_for_ncu.err
echo "rc $?" >> gpurun_out/r02c_bench_for_ncu.err
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02c_ncu_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --atlas-cells 0 --candidate-piles 0 > gpurun_out/r02c_ncu_launches.log 2>&1
echo "rc $?" >> gpurun_out/r02c_ncu_launches.log
timeout 120 python scripts/time_partitions.py --cells 4000 --no-cpu --repeat 1 > gpurun_out/r02c_tp4000.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:optimize_partitions_kernel -c 1 -o gpurun_out/r02c_ncu_optimize_4000 python scripts/time_partitions.py --cells 4000 --no-cpu --repeat 1 > gpurun_out/r02c_ncu_optimize.log 2>&1
echo "rc $?" >> gpurun_out/r02c_ncu_optimize.log
