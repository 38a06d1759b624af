#!/bin/bash
# round 2, GPU session 10 (one GPU): differential fuzz of the code generator on the persistent kernel; final ncu of the Equihash rounds
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
timeout 900 python tests/fuzz_periods.py --periods 400 --seed 3 --nonces 12 > gpurun_out/s10_fuzz_e0.json 2> gpurun_out/s10_fuzz.err
timeout 600 python tests/fuzz_periods.py --periods 200 --seed 4 --epoch 448 --nonces 12 > gpurun_out/s10_fuzz_e448.json 2>> gpurun_out/s10_fuzz.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"eq_collide|eq_generate" -s 30 -c 10 -o gpurun_out/s10_equihash python tools/bench_equihash.py --nonces 2 > gpurun_out/s10_ncu.log 2>&1
cat gpurun_out/s10_fuzz_e0.json gpurun_out/s10_fuzz_e448.json; tail -n 3 gpurun_out/s10_fuzz.err
