set -x
mkdir -p gpurun_out
timeout 300 python scripts/time_upload.py > gpurun_out/r2aj_upload.log 2>&1
echo "rc $?" >> gpurun_out/r2aj_upload.log
MCB_BENCH_TRACE=1 timeout 900 python bench.py --steps 4 --warmup 3 --atlas-cells 0 --no-cpu-baseline --candidate-piles 0 > gpurun_out/r2aj_bench.json 2> gpurun_out/r2aj_bench.err
echo "rc $?" >> gpurun_out/r2aj_bench.err
