# dev helper: run the full bench several times, each bounded, to catch rare hangs (prints one short line per run)
for i in 1 2 3 4; do
  SGM_B200_BENCH_WATCHDOG_S=200 timeout 260 python bench.py > gpurun_out/soak_$i.json 2> gpurun_out/soak_$i.err
  echo "run $i rc=$? $(python -c "
import json
try:
    d=json.loads(open('gpurun_out/soak_$i.json').read().strip().splitlines()[-1]); print(round(d['value'],1), round(d['e2e']['value'],1), round(d['stage_ms_per_frame']['aggregate'],3), d['cpu_baseline'] and d['cpu_baseline']['value'])
except Exception as e: print('no line', e)")"
  tail -c 600 gpurun_out/soak_$i.err
done
