mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_stream_gpu.py -m gpu -x -q 2>&1 | tail -n 4
timeout 120 python tests/stream_config4.py 1 4096 5 | tee gpurun_out/r02_v23_config4_n1.json
