# dev helper for one `gpurun --gpus 8` call: the frame-stream scaling lines of bench.py at 2 / 4 / 8 GPUs (torchrun, one
# rank per GPU, independent pairs, no collective) and the integer mode at 8.  usage: bash tests/scale_run.sh <tag>
tag=${1:-vX}
out=gpurun_out
mkdir -p $out
python bench.py --no-cpu-baseline > $out/r01_${tag}_bench_n1.json 2> $out/scale.err
for n in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
      bench.py --gpus $n --no-cpu-baseline > $out/r01_${tag}_bench_n$n.json 2>> $out/scale.err
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 \
    bench.py --gpus 8 --mode integer --no-cpu-baseline > $out/r01_${tag}_bench_int_n8.json 2>> $out/scale.err
for f in $out/r01_${tag}_bench_n*.json $out/r01_${tag}_bench_int_n8.json; do
  python -c "
import json,sys
d=json.loads(open('$f').read().strip().splitlines()[-1]); print('$f', d['n_gpus'], round(d['value'],1), round(d['e2e']['value'],1), d['clocks']['sm_mhz'], d['clocks']['reasons'])"
done
nproc
