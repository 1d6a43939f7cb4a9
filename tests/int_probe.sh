# dev helper: parity + both bench modes (one gpurun call)
python -m pytest tests -m gpu -x -q 2>&1 | tail -1
for m in reference integer; do
python bench.py --mode $m --no-cpu-baseline --steps 10 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$m', round(d['value'],1), round(d['e2e']['value'],1), {k:round(v,3) for k,v in d['stage_ms_per_frame'].items()})"
done
