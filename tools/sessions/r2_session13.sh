#!/bin/bash
# round 2, GPU session 13 (one GPU): the default bench command (120 steps = 10 s of steady state) on the final code
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/s13_bench_n1.json 2> gpurun_out/s13_bench_n1.err
echo "bench rc=$?"; cut -c1-700 gpurun_out/s13_bench_n1.json
