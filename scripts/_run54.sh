set -x
mkdir -p gpurun_out
MCB_NO_SMALL_UPLOADS=1 timeout 150 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --atlas-cells 60000 --candidate-piles 0 > gpurun_out/r02c_bench_guard.json 2> gpurun_out/r02c_bench_guard.err
echo "rc $?" >> gpurun_out/r02c_bench_guard.err
