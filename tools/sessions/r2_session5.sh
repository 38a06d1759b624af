#!/bin/bash
# round 2, GPU session 5 (one GPU): Equihash engine with 16-byte row accesses, full GPU suite
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_equihash.py -x -q -m gpu --timeout 300 > gpurun_out/s5_pytest_eq.log 2>&1
echo "pytest equihash rc=$?" > gpurun_out/s5_status.txt
timeout 300 python tools/bench_equihash.py --nonces 64 > gpurun_out/s5_equihash.json 2> gpurun_out/s5_equihash.err
echo "bench equihash rc=$?" >> gpurun_out/s5_status.txt
timeout 1200 python -m pytest tests -x -q -m gpu --timeout 400 > gpurun_out/s5_pytest_all.log 2>&1
echo "pytest all rc=$?" >> gpurun_out/s5_status.txt
timeout 300 ncu --set full --clock-control none --import-source on -k regex:eq_collide -s 12 -c 1 -o gpurun_out/s5_equihash python tools/bench_equihash.py --nonces 2 > gpurun_out/s5_ncu.log 2>&1
cat gpurun_out/s5_status.txt; tail -n 4 gpurun_out/s5_pytest_eq.log; tail -n 4 gpurun_out/s5_pytest_all.log; cat gpurun_out/s5_equihash.json
