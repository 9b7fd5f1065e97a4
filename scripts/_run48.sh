set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_select.py tests/test_gpu_bench_pile.py tests/test_gpu_seeds.py tests/test_gpu_atlas.py -x -q -m gpu > gpurun_out/r2am_tests.log 2>&1
echo "rc $?" >> gpurun_out/r2am_tests.log
MCB_BENCH_TRACE=1 timeout 900 python bench.py --steps 6 --warmup 3 --atlas-cells 0 --no-cpu-baseline --candidate-piles 0 > gpurun_out/r2am_bench.json 2> gpurun_out/r2am_bench.err
echo "rc $?" >> gpurun_out/r2am_bench.err
