mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q > gpurun_out/r02_call7_pytest.log 2>&1; echo "suite rc=$?"; tail -5 gpurun_out/r02_call7_pytest.log
for cfg in "sgm 375 1242 128" "bmsad 375 1242 64" "bm 375 1242 64" "sgm 1080 1920 256"; do
  set -- $cfg
  python tests/run_one.py $1 $2 $3 $4 6 > gpurun_out/run_one_$1_$4.log 2>&1 || { echo "run_one $cfg failed"; tail -3 gpurun_out/run_one_$1_$4.log; continue; }
  tail -1 gpurun_out/run_one_$1_$4.log
  ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_v22_launches_$1_d$4.csv python tests/run_one.py $1 $2 $3 $4 6 > gpurun_out/ncu_one.log 2>&1
  python - "$1" "$4" <<'PY'
import csv,sys,collections
rows=[r for r in csv.reader(open(f'gpurun_out/r02_v22_launches_{sys.argv[1]}_d{sys.argv[2]}.csv')) if len(r)>5 and r[0].isdigit()]
agg=collections.OrderedDict()
for r in rows: agg.setdefault(r[4].split('(')[0][-60:],[]).append(float(r[-1]))
print(sys.argv[1:], "launches", len(rows))
for k,v in agg.items():
    if 'median' in k or 'classify' in k or 'speckle' in k: print("  %-62s n=%2d  last %8.1f us" % (k,len(v),v[-1]/1e3))
PY
done
