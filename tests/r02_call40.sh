# Median by rank counting instead of bisection: parity, then phase timers (tuning build) and post-stage times (product).
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "median or kitti_size or bm_ or stages_bit_exact or full_size or fuzz" 2>&1 | tail -n 2
for cfg in "sgm 375 1242 128" "bmsad 375 1242 64" "bm 375 1242 64" "sgm 1080 1920 256"; do
  echo "== $cfg"; timeout 120 python tests/run_one.py $cfg 6 2>&1 | tail -n 1
  SGM_B200_LIBRARY=$PWD/stereo_matching_b200/libsgm_b200_tuning.so timeout 120 python tests/run_one.py $cfg 3 2>&1 | grep median_solve | tail -n 1
done
