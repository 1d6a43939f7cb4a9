mkdir -p gpurun_out
bash tests/collect_profiles.sh v23 reference
BENCH_EXTRA="--no-configs" bash tests/collect_profiles.sh v23int integer
ls -la gpurun_out | grep v23
