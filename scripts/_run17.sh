set -x
mkdir -p gpurun_out
timeout 1700 python scripts/time_candidates.py --cpu --out gpurun_out/r2q_candidates_16000.json > gpurun_out/r2q_tc16000.log 2>&1
echo "rc $?" >> gpurun_out/r2q_tc16000.log
