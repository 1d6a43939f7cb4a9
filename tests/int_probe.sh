# dev helper: median skew experiments
for coop in 241 243 244; do
SGM_B200_MEDIAN_COOP=$coop python -m pytest tests -m gpu -x -q -k "stages_bit_exact or kitti_size" 2>&1 | tail -1
for m in reference integer; do
SGM_B200_MEDIAN_COOP=$coop python bench.py --mode $m --no-cpu-baseline --steps 10 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$m coop $coop', round(d['value'],1), round(d['e2e']['value'],1), {k:round(v,3) for k,v in d['stage_ms_per_frame'].items()})"
done
done
