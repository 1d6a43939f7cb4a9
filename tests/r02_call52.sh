mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 150 python tests/stream_config4.py 8 4096 5 | tee gpurun_out/r02_v30_config4_n8.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --steps 20 --warmup 3 --no-configs --no-cpu-baseline > gpurun_out/r02_v30_bench_n8.json 2> gpurun_out/r02_v30_bench_n8.err; echo "n8 rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r02_v30_bench_n8.json') if l.startswith('{')][-1])
print('N 8 value',round(d['value'],1),'e2e',round(d['e2e']['value'],1),'n_gpus',d['n_gpus'],d['clocks'])
PY
