#!/bin/bash
# round 2, GPU session 2 (one GPU): full GPU test suite, LDS microbenchmark, K7 in-flight knob, bench line with paths, ncu
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu --timeout 400 --durations=12 -s > gpurun_out/s2_pytest_all.log 2>&1
echo "pytest all rc=$?" > gpurun_out/s2_status.txt
timeout 120 ./build/nx_ubench lds > gpurun_out/s2_ubench_lds.json 2> gpurun_out/s2_ubench.err
timeout 120 ./build/nx_ubench int >> gpurun_out/s2_ubench_lds.json 2>> gpurun_out/s2_ubench.err
for q in 4 8; do NX_ETHASH_INFLIGHT=$q timeout 200 python - <<PY >> gpurun_out/s2_k7.txt 2>&1
import sys; sys.path.insert(0, ".")
import nodexa_chain_core_b200 as nx
from nodexa_chain_core_b200 import ethash
ctx = nx.EpochContext(448, 0)
hd = ethash.keccak256(b"paths")
ethash.hash_batch(ctx, hd, 0, 1 << 12)
best = 1e9
for r in range(3):
    ethash.hash_batch(ctx, hd, (r + 1) << 32, 1 << 21)
    best = min(best, ctx.last_timing()[0])
print("K7 inflight $q:", (1 << 21) / best / 1e3, "MH/s")
PY
done
timeout 600 python bench.py > gpurun_out/s2_bench_n1.json 2> gpurun_out/s2_bench_n1.err
echo "bench rc=$?" >> gpurun_out/s2_status.txt
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/s2_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-paths > gpurun_out/s2_ncu_launches.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:nx_kawpow_kernel -s 3 -c 1 -o gpurun_out/s2_kawpow python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-paths > gpurun_out/s2_ncu_full.log 2>&1
cat gpurun_out/s2_status.txt; tail -n 25 gpurun_out/s2_pytest_all.log; cat gpurun_out/s2_k7.txt; cat gpurun_out/s2_ubench_lds.json; cut -c1-1500 gpurun_out/s2_bench_n1.json
