# dev helper for one gpurun call: GPU parity suite, headline bench line, every BASELINE config (not a pytest)
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py --no-cpu-baseline --steps 10 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['e2e']['value'],1), {k:round(v,3) for k,v in d['stage_ms_per_frame'].items()})"
python tests/perf_configs.py --frames 60 | python -c "
import json,sys
for ln in sys.stdin:
    d=json.loads(ln); print(d['config'][:40], round(d['frames_per_s'],1), d['stage_ms'])"
