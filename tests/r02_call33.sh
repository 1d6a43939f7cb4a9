# Product library: full GPU suite and the bench (all configs).
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r02_call33_pytest.log 2>&1; echo "suite rc=$?"; tail -n 3 gpurun_out/r02_call33_pytest.log
timeout 600 python bench.py > gpurun_out/r02_v28_bench_line.json 2> gpurun_out/r02_v28_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
j=json.loads(open('gpurun_out/r02_v28_bench_line.json').read())
print(j['value'], j['e2e']['value'], j['roofline']['frac'], j['stage_ms_per_frame'], j['clocks'])
for c in j['configs']: print(c['config'][:60], round(c['frames_per_s_e2e'],1), c['stage_ms_per_frame'], c.get('roofline',{}).get('frac'))
print(j['blocking'])
PY
