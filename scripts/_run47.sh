set -x
mkdir -p gpurun_out
MCB_BENCH_PROFILE_E2E=1 timeout 900 python bench.py --steps 6 --warmup 3 --atlas-cells 0 --no-cpu-baseline --candidate-piles 0 > gpurun_out/r2al_bench.json 2> gpurun_out/r2al_bench.err
echo "rc $?" >> gpurun_out/r2al_bench.err
