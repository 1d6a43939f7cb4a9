# (NW, HWW) shapes under the scheduler-placement rule, lone-frame aggregate time and throughput with 5 frames in flight.
export SGM_B200_LIBRARY=$PWD/stereo_matching_b200/libsgm_b200_tuning.so
for rep in 1 2; do for cfg in 0 19 20 21 22 23 24 25; do
  printf "cfg %2d: " $cfg
  SGM_B200_SWEEP_CFG=$cfg timeout 120 python tests/run_one.py sgm 375 1242 128 12 2>&1 | grep -v median_solve | tail -n 1 | sed "s/.*'aggregate': \([0-9.]*\).*/\1/"
done; done
for cfg in 0 20 21 22; do
  printf "bench cfg %2d: " $cfg
  SGM_B200_SWEEP_CFG=$cfg timeout 300 python bench.py --no-configs --no-cpu-baseline 2>/dev/null | python -c "import json,sys; j=json.loads(sys.stdin.read()); print(j['value'], j['e2e']['value'], j['stage_ms_per_frame']['aggregate'], j['clocks'])"
done
for icfg in 0 7 8 9; do
  printf "integer cfg %2d: " $icfg
  SGM_B200_ISWEEP_CFG=$icfg timeout 120 python tests/run_one.py sgm 375 1242 128 12 integer 2>&1 | grep -v median_solve | tail -n 1 | sed "s/.*'aggregate': \([0-9.]*\).*/\1/"
done
