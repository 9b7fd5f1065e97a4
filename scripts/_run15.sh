set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_candidates.py -x -q -m gpu > gpurun_out/r2o_t_cand.log 2>&1
echo "rc $?" >> gpurun_out/r2o_t_cand.log
