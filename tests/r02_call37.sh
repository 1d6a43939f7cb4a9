timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "kitti_size or ragged or stages_bit_exact" 2>&1 | tail -n 2
for i in 1 2; do timeout 120 python tests/run_one.py sgm 375 1242 128 12 2>&1 | tail -n 1; done
timeout 120 python tests/run_one.py sgm 1080 1920 256 6 2>&1 | tail -n 1
