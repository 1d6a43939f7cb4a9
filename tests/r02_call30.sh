# Where do the block-edge warps find the neighbour block's vector? (tuning-build debug bit 128 prints counters per warp)
export SGM_B200_LIBRARY=$PWD/stereo_matching_b200/libsgm_b200_tuning.so
SGM_B200_SWEEP_DEBUG=128 timeout 120 python tests/run_one.py sgm 375 1242 128 3 2>&1 | grep -v median_solve | tail -n 20
SGM_B200_SWEEP_DEBUG=128 timeout 120 python tests/run_one.py sgm 375 1242 64 3 2>&1 | grep -v median_solve | tail -n 9
