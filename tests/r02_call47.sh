# Nap of the waiting warps in the median solver (0 / 100 ns), effect on throughput with frames in flight (tuning build, same box).
export SGM_B200_LIBRARY=$PWD/stereo_matching_b200/libsgm_b200_tuning.so
for rep in 1 2; do for nap in 0 100; do
  printf "nap %3d: " $nap
  SGM_B200_MEDIAN_NAP=$nap timeout 400 python bench.py --no-cpu-baseline 2>gpurun_out/nap_$nap.err | grep '^{"metric' | python -c "
import json,sys
j=json.loads(sys.stdin.read())
print(round(j['value'],1), round(j['e2e']['value'],1), [round(c['frames_per_s_e2e'],1) for c in j['configs']])"
done; done
