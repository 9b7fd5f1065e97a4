set -x
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2v_topo.txt 2>&1
lscpu | head -n 30 > gpurun_out/r2v_lscpu.txt 2>&1
for d in /sys/bus/pci/devices/*; do if [ -f $d/numa_node ] && grep -q 0x10de $d/vendor 2>/dev/null; then echo $d $(cat $d/numa_node) $(cat $d/class); fi; done > gpurun_out/r2v_numa.txt 2>&1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 8 --steps 10 --warmup 3 --atlas-cells 0 --oversized-cells 0 > gpurun_out/r2v_bench_n8.json 2> gpurun_out/r2v_bench_n8.err
echo "rc $?" >> gpurun_out/r2v_bench_n8.err
