# Last word of the round: exactly what the driver runs, on the final tree.
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 2
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r02_call56_pytest.log 2>&1; echo "suite rc=$?"; tail -n 2 gpurun_out/r02_call56_pytest.log
timeout 500 python bench.py > gpurun_out/r02_final_bench_line.json 2> gpurun_out/r02_final_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02_final_bench_line.json'))
print('value',round(d['value'],1),'e2e',round(d['e2e']['value'],1),d['stage_ms_per_frame'],'frac',round(d['roofline']['frac'],4),'inflight',d['inflight_frames_per_gpu'])
print(d['clocks'], d['cpu_baseline']['value'], d['blocking']['blocking_fps'])
for c in d['configs']: print(c['config'][:60], round(c['frames_per_s_e2e'],1))
PY
