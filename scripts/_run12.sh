set -x
mkdir -p gpurun_out
timeout 1700 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --candidate-piles 0 --atlas-cells 500000 --atlas-extra 2000000 > gpurun_out/r2l_bench_c3c4.json 2> gpurun_out/r2l_bench_c3c4.err
echo "rc $?" >> gpurun_out/r2l_bench_c3c4.err
free -g >> gpurun_out/r2l_bench_c3c4.err
