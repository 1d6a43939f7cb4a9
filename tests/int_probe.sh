# dev helper: integer-mode timing experiments (one gpurun call)
python -m pytest tests -m gpu -x -q -k "integer" 2>&1 | tail -3
for cfg in 0 1 2 4 5 6; do
SGM_B200_ISWEEP_CFG=$cfg python bench.py --mode integer --inflight 5 --no-cpu-baseline --steps 10 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('integer cfg $cfg', round(d['value'],1), round(d['e2e']['value'],1), {k:round(v,3) for k,v in d['stage_ms_per_frame'].items()}, round(d['roofline']['frac'],3))"
done
