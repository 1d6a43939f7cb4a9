mkdir -p gpurun_out
timeout 700 python -m pytest tests -m gpu -x -q > gpurun_out/r02_call14_pytest.log 2>&1; echo "suite rc=$?"; tail -n 6 gpurun_out/r02_call14_pytest.log
BENCH_EXTRA="" bash tests/collect_profiles.sh v23int integer
ls -la gpurun_out | grep v23int
