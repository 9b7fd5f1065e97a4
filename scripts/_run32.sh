set -x
mkdir -p gpurun_out
timeout 3000 python -m pytest tests -x -q -m gpu --durations=15 > gpurun_out/r2_final_gpu_tests.log 2>&1
echo "rc $?" >> gpurun_out/r2_final_gpu_tests.log
timeout 1500 python bench.py > gpurun_out/r2_final_bench.json 2> gpurun_out/r2_final_bench.err
echo "rc $?" >> gpurun_out/r2_final_bench.err
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2_final_smoke.log 2>&1
echo "rc $?" >> gpurun_out/r2_final_smoke.log
