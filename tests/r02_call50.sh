# Columns per block in the 19-column kernel: 16 (78 blocks/view: too many), 17 (73), 18 (69), 19 (66): throughput and lone-frame time.
export SGM_B200_LIBRARY=$PWD/stereo_matching_b200/libsgm_b200_tuning.so
for rep in 1 2; do for cols in 17 18 19; do
  printf "cols %2d: " $cols
  SGM_B200_SWEEP_CFG=50 SGM_B200_SWEEP_COLS=$cols timeout 400 python bench.py --no-cpu-baseline --no-configs 2>gpurun_out/cols_$cols.err | grep '^{"metric' | python -c "
import json,sys
j=json.loads(sys.stdin.read())
print(round(j['value'],1), round(j['e2e']['value'],1), round(j['stage_ms_per_frame']['aggregate'],4))"
done; done
