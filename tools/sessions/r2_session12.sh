#!/bin/bash
# round 2, GPU session 12 (one GPU): lifecycle soak on the persistent kernel (linked blocks across periods and an epoch boundary)
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
timeout 900 python tools/soak_chain.py --blocks 150 --first-height 7425 > gpurun_out/s12_soak.json 2> gpurun_out/s12_soak.err
echo "soak rc=$?"; cut -c1-1500 gpurun_out/s12_soak.json; tail -n 3 gpurun_out/s12_soak.err
