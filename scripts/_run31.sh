set -x
mkdir -p gpurun_out
timeout 1500 ncu --set full --clock-control none --import-source on -k 'regex:cell_gaps' -c 2 -o gpurun_out/r02b_ncu_gaps python scripts/time_candidates.py --cells 8000 > gpurun_out/r02b_ncu_gaps.log 2>&1
echo "rc $?" >> gpurun_out/r02b_ncu_gaps.log
