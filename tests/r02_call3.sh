mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r02_call3_pytest.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r02_call3_pytest.log
bash tests/collect_profiles.sh v17
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02_v17_bench_line.json'))
print(round(d['value'],1), round(d['e2e']['value'],1), d['stage_ms_per_frame'], d['roofline']['frac'])
PY
