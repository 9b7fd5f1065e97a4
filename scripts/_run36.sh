set -x
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N > gpurun_out/r2_final_bench_n$N.json 2> gpurun_out/r2_final_bench_n$N.err
echo "rc $?" >> gpurun_out/r2_final_bench_n$N.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 -m pytest tests/test_gpu_oversized_nccl.py -x -q -m gpu > gpurun_out/r2_final_nccl_test_n$N.log 2>&1
echo "rc $?" >> gpurun_out/r2_final_nccl_test_n$N.log
