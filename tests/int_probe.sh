# dev helper: integer-mode timing experiments (one gpurun call)
python -m pytest tests -m gpu -x -q -k "integer" 2>&1 | tail -3
for cfg in "integer 3" "integer 4" "integer 6" "reference 3" "reference 4" "reference 5"; do
set -- $cfg
python bench.py --mode $1 --inflight $2 --no-cpu-baseline --steps 10 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$1 inflight $2', round(d['value'],1), round(d['e2e']['value'],1), {k:round(v,3) for k,v in d['stage_ms_per_frame'].items()}, round(d['roofline']['frac'],3))"
done
