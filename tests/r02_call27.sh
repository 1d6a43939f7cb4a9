# Optimistic three-path step of the edge warps (tuning-build debug bit 64 = off): parity, then A/B.
export SGM_B200_LIBRARY=$PWD/stereo_matching_b200/libsgm_b200_tuning.so
timeout 400 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "kitti_size or ragged or stages_bit_exact or fuzz or integer or wide_rows" 2>&1 | tail -n 3
agg() { grep -v median_solve | tail -n 1 | sed "s/.*'aggregate': \([0-9.]*\).*/\1/"; }
for rep in 1 2 3; do for dbg in 0 64; do
  printf "debug %2d D=128: " $dbg; SGM_B200_SWEEP_DEBUG=$dbg timeout 120 python tests/run_one.py sgm 375 1242 128 12 2>&1 | agg
done; done
for dbg in 0 64; do
  printf "debug %2d D=64: " $dbg; SGM_B200_SWEEP_DEBUG=$dbg timeout 120 python tests/run_one.py sgm 375 1242 64 12 2>&1 | agg
  printf "debug %2d integer D=128: " $dbg; SGM_B200_SWEEP_DEBUG=$dbg timeout 120 python tests/run_one.py sgm 375 1242 128 12 integer 2>&1 | agg
  printf "debug %2d 1080p D=256: " $dbg; SGM_B200_SWEEP_DEBUG=$dbg timeout 120 python tests/run_one.py sgm 1080 1920 256 6 2>&1 | agg
done
