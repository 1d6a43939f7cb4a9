# With 146 sweep blocks (two SMs free), is the median still on the critical path of the pipeline?  Slow it down on purpose and look.
export SGM_B200_LIBRARY=$PWD/stereo_matching_b200/libsgm_b200_tuning.so
for small in 0 128 0 128; do
  printf "median threads %3d: " $small
  SGM_B200_MEDIAN_SMALL=$small timeout 300 python bench.py --no-cpu-baseline --no-configs 2>/dev/null | grep '^{"metric' | python -c "
import json,sys
j=json.loads(sys.stdin.read())
print(round(j['value'],1), round(j['e2e']['value'],1), round(j['stage_ms_per_frame']['post'],3))"
done
