#!/bin/bash
# round 2, GPU session 7 (one GPU): Equihash engine after each change - parity tests, Sol/s, launch list
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_equihash.py -x -q -m gpu --timeout 300 > gpurun_out/s7_pytest_eq.log 2>&1
echo "pytest equihash rc=$?" > gpurun_out/s7_status.txt
timeout 300 python tools/bench_equihash.py --nonces 64 > gpurun_out/s7_equihash.json 2> gpurun_out/s7_equihash.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -s 40 -c 24 --csv --log-file gpurun_out/s7_eq_launches.csv python tools/bench_equihash.py --nonces 4 > gpurun_out/s7_ncu_l.log 2>&1
cat gpurun_out/s7_status.txt; tail -n 3 gpurun_out/s7_pytest_eq.log; cat gpurun_out/s7_equihash.json
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/s7_eq_launches.csv')) if len(r)>10 and r[0].isdigit()]
for r in rows[:13]:
    print(r[4].split('(')[0][-28:], float(r[-1])/1e3, 'us')
PY
