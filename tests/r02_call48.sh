# Does a long-running single block (the median) hold up the 148-block cooperative sweeps of the other frames in flight?
# Experiment: median block of 128 / 256 threads with a 1024-candidate chunk (fits beside a sweep block) against the 512-thread one.
export SGM_B200_LIBRARY=$PWD/stereo_matching_b200/libsgm_b200_tuning.so
for rep in 1 2; do for small in 0 128 256; do
  printf "median threads %3d: " $small
  SGM_B200_MEDIAN_SMALL=$small timeout 400 python bench.py --no-cpu-baseline --no-configs 2>gpurun_out/small_$small.err | grep '^{"metric' | python -c "
import json,sys
j=json.loads(sys.stdin.read())
print(round(j['value'],1), round(j['e2e']['value'],1), j['stage_ms_per_frame'])"
done; done
