# PF = 2 shapes under the scheduler-placement rule: lone-frame aggregate time (D = 128, D = 64, integer) and bench throughput.
export SGM_B200_LIBRARY=$PWD/stereo_matching_b200/libsgm_b200_tuning.so
agg() { grep -v median_solve | tail -n 1 | sed "s/.*'aggregate': \([0-9.]*\).*/\1/"; }
for rep in 1 2; do for cfg in 0 16 24 26 27 28 29; do
  printf "cfg %2d D=128: " $cfg; SGM_B200_SWEEP_CFG=$cfg timeout 120 python tests/run_one.py sgm 375 1242 128 12 2>&1 | agg
done; done
for cfg in 0 16 21 24 26 28; do
  printf "cfg %2d D=64: " $cfg; SGM_B200_SWEEP_CFG=$cfg timeout 120 python tests/run_one.py sgm 375 1242 64 12 2>&1 | agg
done
for icfg in 0 7 8 9 10 11 12; do
  printf "integer cfg %2d D=128: " $icfg; SGM_B200_ISWEEP_CFG=$icfg timeout 120 python tests/run_one.py sgm 375 1242 128 12 integer 2>&1 | agg
done
for cfg in 0 24 26 16; do
  printf "bench cfg %2d: " $cfg
  SGM_B200_SWEEP_CFG=$cfg timeout 300 python bench.py --no-configs --no-cpu-baseline 2>gpurun_out/bench_cfg$cfg.err | grep '^{"metric' | python -c "import json,sys; j=json.loads(sys.stdin.read()); print(j['value'], j['e2e']['value'], j['stage_ms_per_frame']['aggregate'], j['clocks'])"
done
SGM_B200_SWEEP_CFG=24 timeout 300 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "kitti_size or ragged or stages_bit_exact or fuzz" 2>&1 | tail -n 3
