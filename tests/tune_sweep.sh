python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for c in 0 1 2 3 4; do echo "cfg $c"; SGM_B200_SWEEP_CFG=$c python tests/diag_gpu.py 375 1242 128 1 8 0 sgm 8 | grep -cE "differ"; SGM_B200_SWEEP_CFG=$c python bench.py --no-cpu-baseline --steps 4 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['e2e']['value'],1), {k:round(v,3) for k,v in d['stage_ms_per_frame'].items()})"; done
