# Remainder folded into the last block (146 blocks at KITTI width instead of 148): parity, then throughput A/B (tuning build).
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "wide_rows or kitti_size or ragged or stages_bit_exact or integer or fuzz or full_size" 2>&1 | tail -n 2
export SGM_B200_LIBRARY=$PWD/stereo_matching_b200/libsgm_b200_tuning.so
for rep in 1 2 3; do for cfg in 0 24; do
  printf "cfg %2d (0 = 146 blocks, 24 = 148 blocks): " $cfg
  SGM_B200_SWEEP_CFG=$cfg timeout 400 python bench.py --no-cpu-baseline --no-configs 2>gpurun_out/fold_$cfg.err | grep '^{"metric' | python -c "
import json,sys
j=json.loads(sys.stdin.read())
print(round(j['value'],1), round(j['e2e']['value'],1), round(j['stage_ms_per_frame']['aggregate'],4))"
done; done
