# vfilter with 16 instead of 8 rows of loads per batch: parity subset and the cost stage time.
timeout 300 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "stages_bit_exact or ragged or kitti_size" 2>&1 | tail -n 2
for i in 1 2 3; do timeout 120 python tests/run_one.py sgm 375 1242 128 12 2>&1 | tail -n 1; done
timeout 120 python tests/run_one.py sgm 375 1242 64 12 2>&1 | tail -n 1
timeout 120 python tests/run_one.py sgm 1080 1920 256 6 2>&1 | tail -n 1
