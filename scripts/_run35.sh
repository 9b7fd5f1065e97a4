set -x
mkdir -p gpurun_out
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2ac_smoke.log 2>&1
echo "rc $?" >> gpurun_out/r2ac_smoke.log
