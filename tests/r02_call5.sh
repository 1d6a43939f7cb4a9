mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r02_call5_pytest.log 2>&1; echo "suite rc=$?"; tail -15 gpurun_out/r02_call5_pytest.log
timeout 300 python bench.py > gpurun_out/r02_v19_bench_line.json 2> gpurun_out/r02_v19_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/r02_v19_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02_v19_bench_line.json'))
print('value',round(d['value'],1),'e2e',round(d['e2e']['value'],1), d['stage_ms_per_frame'], 'frac',round(d['roofline']['frac'],4))
print('parity',d['e2e']['parity_first_frames'],'blocking',d.get('blocking'),'cpu',d.get('cpu_baseline'))
for c in d.get('configs',[]): print(c.get('config','')[:60], round(c.get('frames_per_s_e2e',0),1), c.get('stage_ms_per_frame'), round(c.get('roofline',{}).get('frac',0),3), c.get('unavailable'))
PY
