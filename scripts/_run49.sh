set -x
mkdir -p gpurun_out
timeout 900 python bench.py --steps 10 --warmup 3 --atlas-cells 0 --no-cpu-baseline --candidate-piles 0 > gpurun_out/r2an_bench_a.json 2> gpurun_out/r2an_bench_a.err
echo "rc $?" >> gpurun_out/r2an_bench_a.err
MCB_BENCH_UPLOAD_FIRST=1 timeout 900 python bench.py --steps 10 --warmup 3 --atlas-cells 0 --no-cpu-baseline --candidate-piles 0 > gpurun_out/r2an_bench_b.json 2> gpurun_out/r2an_bench_b.err
echo "rc $?" >> gpurun_out/r2an_bench_b.err
