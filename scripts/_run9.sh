set -x
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -x -q -m gpu > gpurun_out/r2i_t_all.log 2>&1
echo "rc $?" >> gpurun_out/r2i_t_all.log
timeout 1200 python bench.py > gpurun_out/r2i_bench.json 2> gpurun_out/r2i_bench.err
echo "rc $?" >> gpurun_out/r2i_bench.err
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2i_bench_ref.json 2> gpurun_out/r2i_bench_ref.err
echo "rc $?" >> gpurun_out/r2i_bench_ref.err
