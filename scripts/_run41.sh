set -x
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29523 bench.py --gpus $N --steps 3 --warmup 3 --oversized-cells 0 > gpurun_out/r2af_bench_n$N.json 2> gpurun_out/r2af_bench_n$N.err
echo "rc $?" >> gpurun_out/r2af_bench_n$N.err
