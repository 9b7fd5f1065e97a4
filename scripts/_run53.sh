set -x
mkdir -p gpurun_out
timeout 240 python bench.py --atlas-pile-steps direct --steps 2 --warmup 3 --no-cpu-baseline --candidate-piles 0 > gpurun_out/r02c_bench_walk_direct.json 2> gpurun_out/r02c_bench_walk_direct.err
echo "rc $?" >> gpurun_out/r02c_bench_walk_direct.err
