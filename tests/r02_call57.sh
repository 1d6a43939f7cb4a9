# ncu launch list of one bench step, final engine (after the same command ran clean without ncu).
timeout 300 python bench.py --steps 1 --no-cpu-baseline --no-configs > gpurun_out/pre_ncu57.log 2>&1; echo "plain rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 45 --csv --log-file gpurun_out/r02_v30_launches_bench_kitti128.csv \
    python bench.py --steps 1 --no-cpu-baseline --no-configs > gpurun_out/ncu_launch57.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv,collections
rows=[r for r in csv.reader(open('gpurun_out/r02_v30_launches_bench_kitti128.csv')) if len(r)>10]
hdr=rows[0]; k=hdr.index('Kernel Name'); v=hdr.index('Metric Value'); u=hdr.index('Metric Unit')
agg=collections.OrderedDict()
for r in rows[1:]:
    t=float(r[v].replace(',',''))*(1e-3 if r[u]=='ns' else 1)
    agg.setdefault(r[k][:60],[]).append(t)
for n,t in agg.items(): print('%-62s n=%2d  median %.1f us'%(n,len(t),sorted(t)[len(t)//2]))
PY
