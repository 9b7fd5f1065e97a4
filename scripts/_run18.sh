set -x
mkdir -p gpurun_out
timeout 600 python scripts/time_candidates.py --out gpurun_out/r2q_candidates_16000_k.json > gpurun_out/r2q_tc16000_k.log 2>&1
echo "rc $?" >> gpurun_out/r2q_tc16000_k.log
