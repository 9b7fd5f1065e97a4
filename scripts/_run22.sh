set -x
mkdir -p gpurun_out
timeout 1500 python bench.py --atlas-cells 0 > gpurun_out/r2u_bench.json 2> gpurun_out/r2u_bench.err
echo "rc $?" >> gpurun_out/r2u_bench.err
