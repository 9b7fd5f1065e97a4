set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_candidates.py tests/test_gpu_deviants.py -x -q -m gpu -k "tiny or corner" > gpurun_out/r2ag_t.log 2>&1
echo "rc $?" >> gpurun_out/r2ag_t.log
