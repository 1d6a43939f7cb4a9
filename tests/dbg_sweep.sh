# dev helper: aggregate stage time with the sweep waits disabled (results are then wrong; timing only)
for cfg in ${CFGS:-0}; do for c in 0 1 2 3; do
SGM_B200_SWEEP_CFG=$cfg SGM_B200_SWEEP_DEBUG=$c python bench.py --no-cpu-baseline --steps 6 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('cfg $cfg debug $c', round(d['value'],1), {k:round(v,3) for k,v in d['stage_ms_per_frame'].items()})"
done; done
