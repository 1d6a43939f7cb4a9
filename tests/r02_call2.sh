# round 2, GPU call 2: the static-scratch entry point alone in a fresh process, then the whole GPU suite (it runs last)
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r02_call2_noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -5 gpurun_out/r02_call2_noalloc_alone.log
python -m pytest tests -m gpu -x -q > gpurun_out/r02_call2_pytest.log 2>&1; echo "suite rc=$?"; tail -8 gpurun_out/r02_call2_pytest.log
python bench.py --no-cpu-baseline --steps 10 > gpurun_out/r02_call2_bench.json 2> gpurun_out/r02_call2_bench.err; echo "bench rc=$?"; head -c 600 gpurun_out/r02_call2_bench.json
