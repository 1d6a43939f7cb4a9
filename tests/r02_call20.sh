export SGM_B200_LIBRARY=$PWD/stereo_matching_b200/libsgm_b200_tuning.so
for rep in 1 2; do for dbg in 0 16; do
  printf "debug %2d (16 = identity mapping): " $dbg
  SGM_B200_SWEEP_DEBUG=$dbg python tests/run_one.py sgm 375 1242 128 12 2>&1 | grep -v median_solve | tail -n 1 | sed "s/.*'aggregate': \([0-9.]*\).*/\1/"
  printf "   D=64: "; SGM_B200_SWEEP_DEBUG=$dbg python tests/run_one.py sgm 375 1242 64 12 2>&1 | grep -v median_solve | tail -n 1 | sed "s/.*'aggregate': \([0-9.]*\).*/\1/"
done; done
timeout 300 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "kitti_size or ragged or stages_bit_exact or fuzz" 2>&1 | tail -n 3
