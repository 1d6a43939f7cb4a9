mkdir -p gpurun_out
python -m pytest tests/test_parity_gpu.py tests/test_stream_gpu.py -m gpu -x -q > gpurun_out/r02_call4_pytest.log 2>&1; echo "suite rc=$?"; tail -6 gpurun_out/r02_call4_pytest.log
bash tests/collect_profiles.sh v18 reference noncu
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02_v18_bench_line.json'))
print(round(d['value'],1), round(d['e2e']['value'],1), d['stage_ms_per_frame'], d['roofline']['frac'])
for l in open('gpurun_out/r02_v18_perf_configs.jsonl'):
    c=json.loads(l); print(c['config'][:58], round(c['frames_per_s'],1), c['stage_ms'])
PY
