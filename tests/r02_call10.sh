mkdir -p gpurun_out
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r02_v23_bench_n2.json 2> gpurun_out/r02_v23_bench_n2.err; echo "n2 rc=$?"; tail -n 3 gpurun_out/r02_v23_bench_n2.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r02_v23_reference_n2.json 2>> gpurun_out/r02_v23_bench_n2.err; echo "ref n2 rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r02_v23_bench_n2.json') if l.startswith('{')][-1])
print('N=2 value',round(d['value'],1),'e2e',round(d['e2e']['value'],1),'n_gpus',d['n_gpus'],d['clocks'])
r=json.loads([l for l in open('gpurun_out/r02_v23_reference_n2.json') if l.startswith('{')][-1]); print('ref', r.get('value'), 'same config', r['config']==d['config'])
PY
