set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2k_gpus.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_oversized_nccl.py -x -q -m gpu > gpurun_out/r2k_t_nccl.log 2>&1
echo "rc $?" >> gpurun_out/r2k_t_nccl.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --atlas-cells 200000 > gpurun_out/r2k_bench_n2_small.json 2> gpurun_out/r2k_bench_n2_small.err
echo "rc $?" >> gpurun_out/r2k_bench_n2_small.err
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 > gpurun_out/r2k_bench_n2.json 2> gpurun_out/r2k_bench_n2.err
echo "rc $?" >> gpurun_out/r2k_bench_n2.err
