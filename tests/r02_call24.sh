# Upper bound for helper warps: edge columns of full blocks without sums / selection / stores (tuning-build debug bit 32, wrong results).
export SGM_B200_LIBRARY=$PWD/stereo_matching_b200/libsgm_b200_tuning.so
agg() { grep -v median_solve | tail -n 1 | sed "s/.*'aggregate': \([0-9.]*\).*/\1/"; }
for rep in 1 2; do for cfg in 0 16 24; do for dbg in 0 32; do
  printf "cfg %2d debug %2d D=128: " $cfg $dbg; SGM_B200_SWEEP_DEBUG=$dbg SGM_B200_SWEEP_CFG=$cfg timeout 120 python tests/run_one.py sgm 375 1242 128 12 2>&1 | agg
done; done; done
for cfg in 0 24; do for dbg in 0 32; do
  printf "cfg %2d debug %2d D=64: " $cfg $dbg; SGM_B200_SWEEP_DEBUG=$dbg SGM_B200_SWEEP_CFG=$cfg timeout 120 python tests/run_one.py sgm 375 1242 64 12 2>&1 | agg
done; done
for icfg in 0 10; do for dbg in 0 32; do
  printf "integer cfg %2d debug %2d D=128: " $icfg $dbg; SGM_B200_SWEEP_DEBUG=$dbg SGM_B200_ISWEEP_CFG=$icfg timeout 120 python tests/run_one.py sgm 375 1242 128 12 integer 2>&1 | agg
done; done
