set -x
mkdir -p gpurun_out
timeout 600 python scripts/time_pile_upload.py > gpurun_out/r2ak_pile_upload.log 2>&1
echo "rc $?" >> gpurun_out/r2ak_pile_upload.log
