# Rows in flight of the down sweep (2 / 4 / 6 / 8) with the up sweep at 2.
export SGM_B200_LIBRARY=$PWD/stereo_matching_b200/libsgm_b200_tuning.so
agg() { grep -v median_solve | tail -n 1 | sed "s/.*'aggregate': \([0-9.]*\).*/\1/"; }
for rep in 1 2 3; do for cfg in 0 40 41 42; do
  printf "cfg %2d D=128: " $cfg; SGM_B200_SWEEP_CFG=$cfg timeout 120 python tests/run_one.py sgm 375 1242 128 12 2>&1 | agg
done; done
SGM_B200_SWEEP_CFG=40 SGM_B200_SWEEP_DEBUG=128 timeout 120 python tests/run_one.py sgm 375 1242 128 2 2>&1 | grep "sweep down" | tail -n 4
