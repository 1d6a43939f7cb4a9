# Frames in flight per GPU with the v30 engine (product build).
for inflight in 3 4 5 6 8; do
  printf "inflight %d: " $inflight
  timeout 300 python bench.py --no-cpu-baseline --no-configs --inflight $inflight 2>gpurun_out/inflight_$inflight.err | grep '^{"metric' | python -c "
import json,sys
j=json.loads(sys.stdin.read())
print(round(j['value'],1), round(j['e2e']['value'],1), j['device_bytes_per_engine'])"
done
