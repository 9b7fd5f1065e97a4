set -x
mkdir -p gpurun_out
timeout 600 python scripts/time_select_tail.py > gpurun_out/r2ah_tail.log 2>&1
echo "rc $?" >> gpurun_out/r2ah_tail.log
