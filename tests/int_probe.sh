# dev helper: row-barrier nap experiment (SGM_B200_SWEEP_DEBUG = nap_ns << 8)
for nap in 0 20 50 100 200 400; do
for m in reference integer; do
SGM_B200_SWEEP_DEBUG=$((nap*256)) python bench.py --mode $m --no-cpu-baseline --steps 10 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$m nap $nap', round(d['value'],1), round(d['e2e']['value'],1), {k:round(v,3) for k,v in d['stage_ms_per_frame'].items()}, round(d['roofline']['frac'],3))"
done
done
