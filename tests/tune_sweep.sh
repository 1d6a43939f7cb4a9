for dbg in 0 1 2 3; do echo "dbg $dbg"; SGM_B200_SWEEP_DEBUG=$dbg python bench.py --no-cpu-baseline --steps 4 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['e2e']['value'],1), {k:round(v,3) for k,v in d['stage_ms_per_frame'].items()})"; done
