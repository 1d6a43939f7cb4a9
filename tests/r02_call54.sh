# The two bench arms as the driver launches them, final engine (v30, 6 frames in flight by default).
timeout 300 python bench.py --impl reference --gpus 1 --steps 20 --warmup 3 > gpurun_out/r02_v30b_reference_arm_line.json 2> gpurun_out/r02_v30b_bench.err; echo "ref rc=$?"
timeout 400 python bench.py --gpus 1 --steps 20 --warmup 3 > gpurun_out/r02_v30b_bench_line.json 2>> gpurun_out/r02_v30b_bench.err; echo "bench rc=$?"; tail -n 3 gpurun_out/r02_v30b_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02_v30b_bench_line.json')); r=json.load(open('gpurun_out/r02_v30b_reference_arm_line.json'))
print('value',round(d['value'],1),'e2e',round(d['e2e']['value'],1),d['stage_ms_per_frame'],'frac',round(d['roofline']['frac'],4),'launches',d['gpu_launches'],'inflight',d['inflight_frames_per_gpu'])
print('ref',round(r['value'],3),'same config',r['config']==d['config'],'ratio e2e',round(d['e2e']['value']/r['value'],1))
print(d['clocks'], d['e2e']['parity_first_frames'])
PY
