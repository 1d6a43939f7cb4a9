mkdir -p gpurun_out
timeout 700 python -m pytest tests -m gpu -x -q > gpurun_out/r02_call9_pytest.log 2>&1; echo "suite rc=$?"; tail -n 4 gpurun_out/r02_call9_pytest.log
timeout 400 python bench.py > gpurun_out/r02_v23_bench_line.json 2> gpurun_out/r02_v23_bench.err; echo "bench rc=$?"; tail -n 3 gpurun_out/r02_v23_bench.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_v23_reference_arm_line.json 2>> gpurun_out/r02_v23_bench.err; echo "ref rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02_v23_bench_line.json'))
print('value',round(d['value'],1),'e2e',round(d['e2e']['value'],1), d['stage_ms_per_frame'], 'frac',round(d['roofline']['frac'],4), 'wall e2e ms', round(d['e2e']['wall_ms_rank0'],1), 'ms_e2e-derived', round(d['steps']*d['frames_per_step']/d['e2e']['value']*1e3,1))
print('blocking',d.get('blocking'))
for c in d.get('configs',[]): print(c.get('config','')[:60], round(c.get('frames_per_s_e2e',0),1), c.get('stage_ms_per_frame'), round(c.get('roofline',{}).get('frac',0),3), c.get('unavailable'))
r=json.load(open('gpurun_out/r02_v23_reference_arm_line.json')); print('reference arm', r.get('value'), r.get('cpu_baseline'), 'same config:', r.get('config')==d['config'])
PY
