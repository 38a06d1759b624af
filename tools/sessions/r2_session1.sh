#!/bin/bash
# round 2, GPU session 1: parity of the persistent kernel, then ceilings and micro-benchmarks
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/s1_gpu.txt
timeout 900 python -m pytest tests/test_gpu_kawpow.py tests/test_gpu_miner.py tests/test_gpu_verify.py -x -q -m gpu > gpurun_out/s1_pytest_kawpow.log 2>&1
echo "pytest kawpow rc=$?" >> gpurun_out/s1_gpu.txt
export SWEEP_STEPS=40
timeout 600 python tools/sweep_tune.py "" "persist=0" "nodag=1" "nomath=1" "threads=512,minblocks=2" "copies=4,threads=1024" > gpurun_out/s1_sweep.txt 2>&1
timeout 300 ./build/nx_ubench all > gpurun_out/s1_ubench.json 2> gpurun_out/s1_ubench.err
timeout 300 python tools/bench_search_sizes.py > gpurun_out/s1_search_sizes.json 2> gpurun_out/s1_search_sizes.err
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/s1_pytest_all.log 2>&1
echo "pytest all rc=$?" >> gpurun_out/s1_gpu.txt
tail -3 gpurun_out/s1_pytest_kawpow.log gpurun_out/s1_pytest_all.log; cat gpurun_out/s1_sweep.txt; cat gpurun_out/s1_ubench.json
