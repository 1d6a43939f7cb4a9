export SGM_B200_LIBRARY=$PWD/stereo_matching_b200/libsgm_b200_tuning.so
for dbg in 0 4 8 12 2 3; do
  echo "== SGM_B200_SWEEP_DEBUG=$dbg"
  SGM_B200_SWEEP_DEBUG=$dbg python tests/run_one.py sgm 375 1242 128 8 2>&1 | grep -v median_solve | tail -n 1
  SGM_B200_SWEEP_DEBUG=$dbg python tests/run_one.py sgm 375 1242 64 8 2>&1 | grep -v median_solve | tail -n 1
done
