set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_atlas.py -x -q -m gpu > gpurun_out/r2r_t_atlas.log 2>&1
echo "rc $?" >> gpurun_out/r2r_t_atlas.log
timeout 1500 python bench.py > gpurun_out/r2r_bench.json 2> gpurun_out/r2r_bench.err
echo "rc $?" >> gpurun_out/r2r_bench.err
