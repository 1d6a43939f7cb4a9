# Product library with the edge-column placement: full GPU suite and the bench; then spare-hardware-warp configs of the tuning build.
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02_call21_pytest.log 2>&1; echo "suite rc=$?"; tail -n 3 gpurun_out/r02_call21_pytest.log
timeout 600 python bench.py > gpurun_out/r02_v25_bench_line.json 2> gpurun_out/r02_v25_bench.err; echo "bench rc=$?"; cat gpurun_out/r02_v25_bench_line.json
export SGM_B200_LIBRARY=$PWD/stereo_matching_b200/libsgm_b200_tuning.so
for rep in 1 2; do for cfg in 0 19 20 21 22; do
  printf "cfg %2d: " $cfg
  SGM_B200_SWEEP_CFG=$cfg timeout 120 python tests/run_one.py sgm 375 1242 128 12 2>&1 | grep -v median_solve | tail -n 1 | sed "s/.*'aggregate': \([0-9.]*\).*/\1/"
done; done
