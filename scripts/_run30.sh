set -x
mkdir -p gpurun_out
timeout 600 python scripts/resident_steps.py --steps 3 > gpurun_out/r02b_steps_plain.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k 'regex:balance_|row_cuts|mirror_prune|sliding_median' --launch-skip 18 -c 9 -o gpurun_out/r02b_ncu_step python scripts/resident_steps.py --steps 3 > gpurun_out/r02b_ncu_step.log 2>&1
echo "rc $?" >> gpurun_out/r02b_ncu_step.log
timeout 900 python scripts/time_candidates.py --cells 8000 > gpurun_out/r02b_tc_plain.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k 'regex:cell_gaps|min_cut|significant_genes|gap_inputs|community_weights' -c 10 -o gpurun_out/r02b_ncu_after python scripts/time_candidates.py --cells 8000 > gpurun_out/r02b_ncu_after.log 2>&1
echo "rc $?" >> gpurun_out/r02b_ncu_after.log
ls -la gpurun_out/*.ncu-rep
