python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 2
timeout 400 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "stages_bit_exact or kitti_size or wide_rows or median" 2>&1 | tail -n 2
timeout 200 python bench.py --no-configs --no-cpu-baseline 2>/dev/null | python -c "import json,sys; j=json.loads(sys.stdin.read()); print(round(j['value'],1), round(j['e2e']['value'],1), j['stage_ms_per_frame'], j['e2e']['parity_first_frames']['ok'])"
