set -x
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 10 --warmup 3 --atlas-cells 0 --no-cpu-baseline --candidate-piles 0 > gpurun_out/r2ao_bench_n2.json 2> gpurun_out/r2ao_bench_n2.err
echo "rc $?" >> gpurun_out/r2ao_bench_n2.err
