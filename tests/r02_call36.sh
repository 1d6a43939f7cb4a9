# vfilter with rolling prefetch: parity subset, stage times, launch list of one frame.
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "kitti_size or ragged or stages_bit_exact or integer_mode_stages or bm_census or fullsize or full_size" 2>&1 | tail -n 3
for i in 1 2; do timeout 120 python tests/run_one.py sgm 375 1242 128 12 2>&1 | tail -n 1; done
timeout 120 python tests/run_one.py sgm 375 1242 64 12 2>&1 | tail -n 1
timeout 120 python tests/run_one.py sgm 1080 1920 256 6 2>&1 | tail -n 1
timeout 120 python tests/run_one.py bm 375 1242 64 12 2>&1 | tail -n 1
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"vfilter|hfilter" -c 12 python tests/run_one.py sgm 375 1242 128 6 2>&1 | grep -E "vfilter|hfilter|gpu__time_duration" | tail -n 8
