#!/bin/bash
# round 2, GPU session 7 (one GPU): Equihash engine after each change - parity tests, Sol/s, launch list
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_equihash.py -x -q -m gpu --timeout 300 > gpurun_out/s7_pytest_eq.log 2>&1
echo "pytest equihash rc=$?" > gpurun_out/s7_status.txt
timeout 300 python tools/bench_equihash.py --nonces 64 > gpurun_out/s7_equihash.json 2> gpurun_out/s7_equihash.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -s 40 -c 24 --csv --log-file gpurun_out/s7_eq_launches.csv python tools/bench_equihash.py --nonces 4 > gpurun_out/s7_ncu_l.log 2>&1
NX_EQ_SWAP=1 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -s 40 -c 24 --csv --log-file gpurun_out/s7_eq_launches_swap.csv python tools/bench_equihash.py --nonces 4 > gpurun_out/s7_ncu_l2.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -s 72 -c 16 --csv --log-file gpurun_out/s7_eq_launches_144.csv python tools/bench_equihash.py --nonces 4 > gpurun_out/s7_ncu_l3.log 2>&1
cat gpurun_out/s7_status.txt; tail -n 3 gpurun_out/s7_pytest_eq.log; cat gpurun_out/s7_equihash.json
python - <<'PY'
import csv
for f in ('s7_eq_launches.csv','s7_eq_launches_swap.csv','s7_eq_launches_144.csv'):
    rows=[r for r in csv.reader(open('gpurun_out/'+f)) if len(r)>10 and r[0].isdigit()]
    print(f, ' '.join('%s:%.0f' % (r[4].split('(')[0].split('::')[-1].replace('eq_','').replace('_kernel',''), float(r[-1])/1e3) for r in rows[:16]))
PY
