# Experiment: sweeps as plain launches instead of cooperative ones (tuning build): throughput with 5 / 3 / 2 frames in flight.
export SGM_B200_LIBRARY=$PWD/stereo_matching_b200/libsgm_b200_tuning.so
for inflight in 5 3; do for plain in 0 1; do
  printf "inflight %d plain %d: " $inflight $plain
  if [ $plain = 1 ]; then export SGM_B200_SWEEP_PLAIN_LAUNCH=1; else unset SGM_B200_SWEEP_PLAIN_LAUNCH; fi
  timeout 200 python bench.py --no-configs --no-cpu-baseline --inflight $inflight 2>gpurun_out/plain_$inflight_$plain.err | grep '^{"metric' | python -c "import json,sys; j=json.loads(sys.stdin.read()); print(j['value'], j['e2e']['value'], j['stage_ms_per_frame']['aggregate'])"
done; done
