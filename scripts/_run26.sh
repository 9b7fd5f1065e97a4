set -x
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 10 --warmup 3 --atlas-cells 0 --oversized-cells 0 > gpurun_out/r2v_bench_n$N.json 2> gpurun_out/r2v_bench_n$N.err
echo "rc $?" >> gpurun_out/r2v_bench_n$N.err
