set -x
mkdir -p gpurun_out
timeout 1500 python bench.py --no-cpu-baseline --candidate-piles 0 --steps 3 --warmup 3 > gpurun_out/r2af_bench.json 2> gpurun_out/r2af_bench.err
echo "rc $?" >> gpurun_out/r2af_bench.err
