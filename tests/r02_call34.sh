mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 150 python tests/stream_config4.py 8 4096 5 | tee gpurun_out/r02_v28_config4_n8.json
timeout 150 python tests/stream_config4.py 4 4096 5 | tee gpurun_out/r02_v28_config4_n4.json
timeout 150 python tests/stream_config4.py 1 4096 5 | tee gpurun_out/r02_v28_config4_n1.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/r02_v28_bench_n8.json 2> gpurun_out/r02_v28_bench_n8.err; echo "n8 rc=$?"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus 4 --steps 20 --warmup 3 --no-configs --no-cpu-baseline > gpurun_out/r02_v28_bench_n4.json 2> gpurun_out/r02_v28_bench_n4.err; echo "n4 rc=$?"
python - <<'PY'
import json
for n in (8,4):
    d=json.loads([l for l in open(f'gpurun_out/r02_v28_bench_n{n}.json') if l.startswith('{')][-1])
    print('N',n,'value',round(d['value'],1),'e2e',round(d['e2e']['value'],1),'n_gpus',d['n_gpus'],d['clocks'])
PY
