export SGM_B200_LIBRARY=$PWD/stereo_matching_b200/libsgm_b200_tuning.so
for cfg in 0 12 14 15 16 17 18 10 11 7 8 4 1 2 3 9; do
  printf "cfg %2d: " $cfg
  SGM_B200_SWEEP_CFG=$cfg python tests/run_one.py sgm 375 1242 128 10 2>&1 | grep -v median_solve | tail -n 1 | sed "s/.*'aggregate': \([0-9.]*\).*/\1/"
done
