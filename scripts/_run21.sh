set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_knn.py tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r2t_t_knn.log 2>&1
echo "rc $?" >> gpurun_out/r2t_t_knn.log
timeout 600 python scripts/resident_steps.py --steps 12 --times > gpurun_out/r2t_steps_bucket.log 2>&1
MCB_RANK_BY_ELEMENT=1 timeout 600 python scripts/resident_steps.py --steps 12 --times > gpurun_out/r2t_steps_element.log 2>&1
