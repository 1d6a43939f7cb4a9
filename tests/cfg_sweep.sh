# dev helper: headline bench under each sweep configuration
for c in 0 20 21 22 23; do
SGM_B200_SWEEP_CFG=$c python bench.py --no-cpu-baseline --steps 10 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('cfg $c', round(d['value'],1), round(d['e2e']['value'],1), {k:round(v,3) for k,v in d['stage_ms_per_frame'].items()})"
done
