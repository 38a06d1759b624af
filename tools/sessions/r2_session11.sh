#!/bin/bash
# round 2, GPU session 11 (one GPU): kernel shapes re-measured on the persistent kernel
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
export SWEEP_STEPS=40
timeout 900 python tools/sweep_tune.py "" "unroll=2" "unroll=4" "unroll=8" "copies=4,threads=512,minblocks=2" "copies=2,threads=256,minblocks=4" "rotimad=1" "daghint=1" > gpurun_out/s11_sweep.txt 2>&1
cat gpurun_out/s11_sweep.txt
