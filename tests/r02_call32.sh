# 13-column shape (1080p, D = 256) and the cost-model shape picker: parity, then A/B against the default shape (tuning build knob).
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "wide_rows or full_size or kitti_size or ragged or stages_bit_exact or integer" 2>&1 | tail -n 3
export SGM_B200_LIBRARY=$PWD/stereo_matching_b200/libsgm_b200_tuning.so
agg() { grep -v median_solve | tail -n 1 | sed "s/.*'aggregate': \([0-9.]*\).*/\1/"; }
for rep in 1 2; do
  printf "1080p D=256 auto (13 columns): "; timeout 120 python tests/run_one.py sgm 1080 1920 256 6 2>&1 | agg
  printf "1080p D=256 default shape:     "; SGM_B200_SWEEP_CFG=13 timeout 120 python tests/run_one.py sgm 1080 1920 256 6 2>&1 | agg
  printf "1080p D=256 integer auto:      "; timeout 120 python tests/run_one.py sgm 1080 1920 256 6 integer 2>&1 | agg
  printf "1080p D=256 integer default:   "; SGM_B200_ISWEEP_CFG=1 timeout 120 python tests/run_one.py sgm 1080 1920 256 6 integer 2>&1 | agg
  printf "1080p D=128 auto:              "; timeout 120 python tests/run_one.py sgm 1080 1920 128 6 2>&1 | agg
  printf "1080p D=128 default:           "; SGM_B200_SWEEP_CFG=13 timeout 120 python tests/run_one.py sgm 1080 1920 128 6 2>&1 | agg
  printf "1242x375 D=256 auto:           "; timeout 120 python tests/run_one.py sgm 375 1242 256 8 2>&1 | agg
  printf "1242x375 D=256 default:        "; SGM_B200_SWEEP_CFG=13 timeout 120 python tests/run_one.py sgm 375 1242 256 8 2>&1 | agg
  printf "760x375 D=128 auto:            "; timeout 120 python tests/run_one.py sgm 375 760 128 8 2>&1 | agg
  printf "760x375 D=128 default:         "; SGM_B200_SWEEP_CFG=13 timeout 120 python tests/run_one.py sgm 375 760 128 8 2>&1 | agg
done
