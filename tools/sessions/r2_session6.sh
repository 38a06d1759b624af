#!/bin/bash
# round 2, GPU session 6 (one GPU): where the Equihash engine's time goes (launch list + full capture of one nonce's rounds)
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
timeout 300 python tools/bench_equihash.py --nonces 64 > gpurun_out/s6_equihash.json 2> gpurun_out/s6_equihash.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -s 40 -c 40 --csv --log-file gpurun_out/s6_eq_launches.csv python tools/bench_equihash.py --nonces 4 > gpurun_out/s6_ncu_l.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"eq_collide|eq_generate" -s 30 -c 10 -o gpurun_out/s6_equihash python tools/bench_equihash.py --nonces 2 > gpurun_out/s6_ncu.log 2>&1
cat gpurun_out/s6_equihash.json; grep -c . gpurun_out/s6_eq_launches.csv
