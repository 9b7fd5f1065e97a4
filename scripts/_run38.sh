set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_atlas.py -x -q -m gpu > gpurun_out/r2ad_t_atlas.log 2>&1
echo "rc $?" >> gpurun_out/r2ad_t_atlas.log
timeout 1500 python bench.py --no-cpu-baseline --candidate-piles 0 --atlas-pile-steps direct --steps 3 --warmup 3 > gpurun_out/r2ad_bench_direct.json 2> gpurun_out/r2ad_bench_direct.err
echo "rc $?" >> gpurun_out/r2ad_bench_direct.err
