#!/bin/bash
# round 2, GPU session 3 (N GPUs, default 2): the multi-process epoch build (one chain + broadcast) and the scaling bench line
N=${1:-2}
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/s3_gpus.txt
if [ "$N" = "2" ]; then
  timeout 900 python -m pytest tests/test_gpu_dag.py -x -q -m gpu -k "two_gpu" --timeout 600 > gpurun_out/s3_pytest_2gpu.log 2>&1
  echo "pytest 2gpu rc=$?" >> gpurun_out/s3_gpus.txt
fi
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 40 --warmup 3 > gpurun_out/s3_bench_n$N.json 2> gpurun_out/s3_bench_n$N.err
echo "bench n$N rc=$?" >> gpurun_out/s3_gpus.txt
cat gpurun_out/s3_gpus.txt; tail -n 5 gpurun_out/s3_pytest_2gpu.log 2>/dev/null; tail -n 5 gpurun_out/s3_bench_n$N.err; cut -c1-3000 gpurun_out/s3_bench_n$N.json
