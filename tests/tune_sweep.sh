python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for leg in 0 1; do
if [ $leg = 1 ]; then export SGM_B200_MEDIAN_LEGACY=1; fi
python bench.py --no-cpu-baseline --steps 6 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['e2e']['value'],1), {k:round(v,3) for k,v in d['stage_ms_per_frame'].items()})"
done
