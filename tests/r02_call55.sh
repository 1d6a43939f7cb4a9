timeout 200 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -n 2
timeout 200 python -m pytest tests/test_stream_gpu.py tests/test_views_gpu.py -m gpu -x -q 2>&1 | tail -n 2
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 2
