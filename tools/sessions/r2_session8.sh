#!/bin/bash
# round 2, GPU session 8 (one GPU): Equihash - which buffer makes every other round slow, and the counter stride
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_equihash.py -x -q -m gpu --timeout 300 > gpurun_out/s8_pytest_eq.log 2>&1
echo "pytest equihash rc=$?" > gpurun_out/s8_status.txt
for st in 1 8 32; do
  NX_EQ_CNT_STRIDE=$st timeout 300 python tools/bench_equihash.py --nonces 64 > gpurun_out/s8_equihash_stride$st.json 2>> gpurun_out/s8_equihash.err
  NX_EQ_CNT_STRIDE=$st timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -s 44 -c 11 --csv --log-file gpurun_out/s8_l_stride$st.csv python tools/bench_equihash.py --nonces 4 > gpurun_out/s8_ncu.log 2>&1
done
for sw in 2 3; do
  NX_EQ_CNT_STRIDE=1 NX_EQ_SWAP=$sw timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -s 44 -c 11 --csv --log-file gpurun_out/s8_l_swap$sw.csv python tools/bench_equihash.py --nonces 4 > gpurun_out/s8_ncu.log 2>&1
done
cat gpurun_out/s8_status.txt; tail -n 3 gpurun_out/s8_pytest_eq.log
python - <<'PY'
import csv, json, glob
for f in sorted(glob.glob('gpurun_out/s8_equihash_stride*.json')):
    d=json.load(open(f)); print(f, [(r['n'], round(r['sol_per_s_device'],1), r['rows_dropped']) for r in d['runs']])
for f in sorted(glob.glob('gpurun_out/s8_l_*.csv')):
    rows=[r for r in csv.reader(open(f)) if len(r)>10 and r[0].isdigit()]
    print(f, ' '.join('%s:%.0f' % (r[4].split('(')[0].split('::')[-1].replace('eq_','').replace('_kernel',''), float(r[-1])/1e3) for r in rows[:11]))
PY
