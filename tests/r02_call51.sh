mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 3
timeout 700 python -m pytest tests -x -q -m gpu > gpurun_out/r02_call51_pytest.log 2>&1; echo "suite rc=$?"; tail -n 4 gpurun_out/r02_call51_pytest.log
timeout 300 python bench.py --impl reference --gpus 1 --steps 20 --warmup 3 > gpurun_out/r02_v30_reference_arm_line.json 2> gpurun_out/r02_v30_bench.err; echo "ref rc=$?"
timeout 400 python bench.py --gpus 1 --steps 20 --warmup 3 > gpurun_out/r02_v30_bench_line.json 2>> gpurun_out/r02_v30_bench.err; echo "bench rc=$?"; tail -n 3 gpurun_out/r02_v30_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02_v30_bench_line.json')); r=json.load(open('gpurun_out/r02_v30_reference_arm_line.json'))
print('value',round(d['value'],1),'e2e',round(d['e2e']['value'],1),d['stage_ms_per_frame'],'frac',round(d['roofline']['frac'],4),'launches',d['gpu_launches'],d['e2e']['gpu_launches'])
print('ref',round(r['value'],3),'same config',r['config']==d['config'],'ratio e2e',round(d['e2e']['value']/r['value'],1), r['cpu_baseline']['child_respawns'])
print(d['clocks'], d['blocking'])
PY
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02_v30_bench_line.json'))
for c in d['configs']: print(c['config'][:60], round(c['frames_per_s_e2e'],1), c['stage_ms_per_frame'], c.get('roofline',{}).get('frac'))
PY
