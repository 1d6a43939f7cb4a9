# Product library (17-column shape + together-step of the edge warps): full GPU suite, bench, launch list; then shapes A/B (tuning).
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r02_call28_pytest.log 2>&1; echo "suite rc=$?"; tail -n 3 gpurun_out/r02_call28_pytest.log
timeout 600 python bench.py > gpurun_out/r02_v27_bench_line.json 2> gpurun_out/r02_v27_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
j=json.loads(open('gpurun_out/r02_v27_bench_line.json').read())
print(j['value'], j['e2e']['value'], j['roofline']['frac'], j['stage_ms_per_frame'], j['clocks'])
for c in j['configs']: print(c['config'][:60], round(c['frames_per_s_e2e'],1), c['stage_ms_per_frame'], c.get('roofline',{}).get('frac'))
print(j['blocking'])
PY
ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 40 --csv --log-file gpurun_out/r02_v27_launches_bench_kitti128.csv \
    python bench.py --steps 1 --no-cpu-baseline --no-configs > gpurun_out/ncu_launch.log 2>&1; echo "ncu rc=$?"
export SGM_B200_LIBRARY=$PWD/stereo_matching_b200/libsgm_b200_tuning.so
agg() { grep -v median_solve | tail -n 1 | sed "s/.*'aggregate': \([0-9.]*\).*/\1/"; }
for rep in 1 2; do for cfg in 0 16 26 27; do
  printf "cfg %2d D=128: " $cfg; SGM_B200_SWEEP_CFG=$cfg timeout 120 python tests/run_one.py sgm 375 1242 128 12 2>&1 | agg
done; done
