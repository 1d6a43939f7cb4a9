for n in 3 5 8 12; do
  printf "inflight %2d: " $n
  python bench.py --inflight $n --no-configs --no-cpu-baseline --steps 20 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['e2e']['value'],1), d['clocks']['sm_mhz'], d['clocks']['reasons'])"
done
