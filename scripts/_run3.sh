set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r2c_noalloc_alone.log 2>&1
echo "rc $?" >> gpurun_out/r2c_noalloc_alone.log
timeout 2400 python -m pytest tests -x -q -m gpu --durations=10 > gpurun_out/r2c_t_all.log 2>&1
echo "rc $?" >> gpurun_out/r2c_t_all.log
timeout 900 python bench.py --no-cpu-baseline --candidate-piles 0 > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err
echo "rc $?" >> gpurun_out/r2c_bench.err
timeout 600 python scripts/time_partitions.py --no-cpu > gpurun_out/r2c_time_partitions.log 2>&1
echo "rc $?" >> gpurun_out/r2c_time_partitions.log
