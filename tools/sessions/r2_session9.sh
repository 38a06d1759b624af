#!/bin/bash
# round 2, GPU session 9 (one GPU): Equihash after the counter/recover changes, then the whole GPU suite and the bench line
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_equihash.py -x -q -m gpu --timeout 300 > gpurun_out/s9_pytest_eq.log 2>&1
echo "pytest equihash rc=$?" > gpurun_out/s9_status.txt
timeout 300 python tools/bench_equihash.py --nonces 64 > gpurun_out/s9_equihash.json 2> gpurun_out/s9_equihash.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -s 48 -c 12 --csv --log-file gpurun_out/s9_eq_launches.csv python tools/bench_equihash.py --nonces 4 > gpurun_out/s9_ncu_l.log 2>&1
timeout 1200 python -m pytest tests -x -q -m gpu --timeout 400 > gpurun_out/s9_pytest_all.log 2>&1
echo "pytest all rc=$?" >> gpurun_out/s9_status.txt
timeout 600 python bench.py > gpurun_out/s9_bench_n1.json 2> gpurun_out/s9_bench_n1.err
echo "bench rc=$?" >> gpurun_out/s9_status.txt
cat gpurun_out/s9_status.txt; tail -n 3 gpurun_out/s9_pytest_eq.log; tail -n 3 gpurun_out/s9_pytest_all.log; cat gpurun_out/s9_equihash.json
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/s9_eq_launches.csv')) if len(r)>10 and r[0].isdigit()]
print(' '.join('%s:%.0f' % (r[4].split('(')[0].split('::')[-1].replace('eq_','').replace('_kernel',''), float(r[-1])/1e3) for r in rows[:12]))
PY
