#!/bin/bash
# round 2, GPU session 7 (one GPU): Equihash engine after each change - parity tests, Sol/s, launch list
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_equihash.py tests/test_gpu_verify.py -x -q -m gpu --timeout 300 > gpurun_out/s7_pytest_eq.log 2>&1
echo "pytest equihash+verify rc=$?" > gpurun_out/s7_status.txt
timeout 300 python tools/bench_equihash.py --nonces 64 > gpurun_out/s7_equihash.json 2> gpurun_out/s7_equihash.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -s 48 -c 12 --csv --log-file gpurun_out/s7_eq_launches.csv python tools/bench_equihash.py --nonces 4 > gpurun_out/s7_ncu_l.log 2>&1
timeout 300 python tools/bench_paths.py --records 65536 > gpurun_out/s7_paths.json 2> gpurun_out/s7_paths.err
cat gpurun_out/s7_status.txt; tail -n 3 gpurun_out/s7_pytest_eq.log; cat gpurun_out/s7_equihash.json
python - <<'PY'
import csv, json
rows=[r for r in csv.reader(open('gpurun_out/s7_eq_launches.csv')) if len(r)>10 and r[0].isdigit()]
print(' '.join('%s:%.0f' % (r[4].split('(')[0].split('::')[-1].replace('eq_','').replace('_kernel',''), float(r[-1])/1e3) for r in rows[:12]))
d=json.load(open('gpurun_out/s7_paths.json')); print({k:d[k] for k in d if k.startswith('K6')})
PY
