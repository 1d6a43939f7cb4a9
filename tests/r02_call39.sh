# ncu --set full of the integer-mode cost / aggregation kernels of one warm frame (v28), after the same command ran clean without ncu.
timeout 300 python bench.py --mode integer --steps 1 --no-cpu-baseline --no-configs > gpurun_out/pre_ncu_int.log 2>&1; echo "plain rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"sgm_sweep|hpair|hamming" -s 12 -c 4 -o gpurun_out/r02_v28int_agg \
    python bench.py --mode integer --steps 1 --no-cpu-baseline --no-configs > gpurun_out/ncu_full_int.log 2>&1; echo "ncu rc=$?"
tail -2 gpurun_out/ncu_full_int.log
