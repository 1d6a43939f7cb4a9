# Helper warps for the block-edge columns: parity first (tuning library, cfg 30), then lone-frame times and throughput.
export SGM_B200_LIBRARY=$PWD/stereo_matching_b200/libsgm_b200_tuning.so
SGM_B200_SWEEP_CFG=30 SGM_B200_ISWEEP_CFG=13 timeout 400 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "kitti_size or ragged or stages_bit_exact or fuzz or integer or fullsize" 2>&1 | tail -n 5
agg() { grep -v median_solve | tail -n 1 | sed "s/.*'aggregate': \([0-9.]*\).*/\1/"; }
for rep in 1 2; do for cfg in 0 24 30 33 34; do
  printf "cfg %2d D=128: " $cfg; SGM_B200_SWEEP_CFG=$cfg timeout 120 python tests/run_one.py sgm 375 1242 128 12 2>&1 | agg
done; done
for cfg in 0 24 30 34; do
  printf "cfg %2d D=64: " $cfg; SGM_B200_SWEEP_CFG=$cfg timeout 120 python tests/run_one.py sgm 375 1242 64 12 2>&1 | agg
done
for icfg in 0 10 13 14; do
  printf "integer cfg %2d D=128: " $icfg; SGM_B200_ISWEEP_CFG=$icfg timeout 120 python tests/run_one.py sgm 375 1242 128 12 integer 2>&1 | agg
done
for cfg in 0 30 16; do
  printf "bench cfg %2d: " $cfg
  SGM_B200_SWEEP_CFG=$cfg timeout 300 python bench.py --no-configs --no-cpu-baseline 2>gpurun_out/bench_cfg$cfg.err | grep '^{"metric' | python -c "import json,sys; j=json.loads(sys.stdin.read()); print(j['value'], j['e2e']['value'], j['stage_ms_per_frame']['aggregate'], j['clocks'])"
done
