#!/bin/bash
# round 2: what the driver runs at round end, on one GPU - smoke(), the GPU suite, both bench arms
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/f_smoke.log 2>&1
echo "smoke rc=$?" > gpurun_out/f_status.txt
timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/f_pytest_gpu.log 2>&1
echo "pytest gpu rc=$?" >> gpurun_out/f_status.txt
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/f_pytest_noalloc_alone.log 2>&1
echo "pytest no-alloc alone rc=$?" >> gpurun_out/f_status.txt
timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/f_bench_ref.json 2> gpurun_out/f_bench_ref.err
echo "bench reference rc=$?" >> gpurun_out/f_status.txt
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 3 > gpurun_out/f_bench_n1_steps20.json 2> gpurun_out/f_bench_n1.err
echo "bench rc=$?" >> gpurun_out/f_status.txt
cat gpurun_out/f_status.txt; tail -n 2 gpurun_out/f_smoke.log; tail -n 3 gpurun_out/f_pytest_gpu.log; cut -c1-400 gpurun_out/f_bench_ref.json; cut -c1-600 gpurun_out/f_bench_n1_steps20.json
